"""CPU suite: the C-ABI library loads, exports every symbol include/transmil_b200.h declares,
and its host-only entry points (index contract, argument validation) behave."""
import ctypes as C
import math
import os
import re

import numpy as np
import pytest

from common import ROOT


@pytest.fixture(scope="module")
def capi(built_lib):
    import transmil_b200 as T
    return T.capi


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "transmil_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tm_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(capi):
    lib = capi.lib()
    names = _declared_symbols()
    assert len(names) >= 25
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/transmil_b200.h but not exported"
    assert set(names) == set(capi.SIGNATURES), "ctypes signature table out of sync with the header"
    assert b"sm_100a" in lib.tm_version()


def test_plan_bit_exact_against_reference_expressions(capi):
    lib = capi.lib()
    p = capi.Plan()
    ns = list(range(1, 20000)) + list(range(20000, 300001, 37)) + [50000, 100000, 200000, 300000]
    for n in ns:
        assert lib.tm_make_plan(n, C.byref(p)) == 0
        G = int(np.ceil(np.sqrt(n)))                                 # TransMIL.py:178
        S = G * G + 1
        padding = 256 - S % 256 if S % 256 > 0 else 0                 # TransMIL.py:190-193
        assert (p.n, p.G, p.N, p.dup, p.S) == (n, G, G * G, G * G - n, S)
        assert (p.l, p.Sp, p.P, p.padding) == (math.ceil(S / 256), S + padding, padding, padding)


def test_index_maps(capi):
    lib = capi.lib()
    for n in (1, 2, 3, 5, 255, 257, 1000, 6000):
        p = capi.make_plan(n)
        src = [lib.tm_dup_source(C.byref(p), t) for t in range(p.N)]
        assert src == list(range(n)) + list(range(p.dup))             # cat([h, h[:, :add]])  TransMIL.py:180
        groups = [lib.tm_landmark_group(C.byref(p), r) for r in range(p.Sp)]
        assert groups == [r // p.l for r in range(p.Sp)] and max(groups) == 255
        assert lib.tm_padded_row(C.byref(p), 0) == p.P                # cls row sits right behind the front pad
        y, x = C.c_int64(), C.c_int64()
        for t in (0, p.N - 1, p.G, min(p.N - 1, p.G + 1)):
            lib.tm_ppeg_pixel(C.byref(p), t, C.byref(y), C.byref(x))
            assert (y.value, x.value) == (t // p.G, t % p.G)


def test_bad_arguments_fail_loudly(capi):
    lib = capi.lib()
    p = capi.Plan()
    assert lib.tm_make_plan(0, C.byref(p)) == -1
    assert b"at least one patch" in lib.tm_last_error()
    with pytest.raises(ValueError):
        capi.make_plan(-5)
    nbytes = C.c_size_t()
    w = capi.Weights()
    assert lib.tm_workspace_bytes(1, 100, C.byref(w), 0, C.byref(nbytes)) != 0   # empty weights struct
    with pytest.raises((ValueError, capi.TransmilError)):
        capi.check(lib.tm_workspace_bytes(1, 100, C.byref(w), 0, C.byref(nbytes)))


def test_workspace_size_host_only(capi):
    lib = capi.lib()
    w = capi.Weights()
    w.fc1_kind, w.in_features, w.n_classes = capi.TM_FC1_SINGLE, 1024, 2
    for f in ("fc1_w0", "fc1_b0", "cls_token", "norm_w", "norm_b", "fc_w", "fc_b"):
        setattr(w, f, 256)   # non-NULL dummies; nothing is dereferenced on the host
    nbytes = C.c_size_t()
    assert lib.tm_workspace_bytes(1, 50000, C.byref(w), 0, C.byref(nbytes)) == 0
    p = capi.make_plan(50000)
    assert nbytes.value > 8 * p.Sp * 512 * 4          # h x3, xn, merged, q, k, v
    assert nbytes.value < 4 << 30


def test_sharded_entry_points_refuse_loudly_without_a_gpu(built_lib):
    """The token-sharded C ABI on a box without a CUDA device: plans are host arithmetic and work; everything that needs the
    device returns a status and a message instead of crashing or pretending (there is no CPU path)."""
    import ctypes as C
    import torch
    import transmil_b200 as T
    L = T.capi.lib()
    sp = T.capi.ShardPlanC()
    assert L.tm_make_shard_plan(200000, 8, 3, C.byref(sp)) == 0
    assert int(sp.world) == 8 and int(sp.rank) == 3 and int(sp.rows) % 256 == 0 and int(sp.ntok) > 0
    assert L.tm_make_shard_plan(200000, 8, 8, C.byref(sp)) != 0 and b"rank" in L.tm_last_error()
    assert L.tm_make_shard_plan(300, 4, 0, C.byref(sp)) != 0            # 2 blocks of 256 rows cannot feed 4 ranks
    assert L.tm_make_shard_plan(0, 1, 0, C.byref(sp)) != 0
    assert L.tm_shard_comm_destroy(None) == 0                            # NULL is a no-op, like free()
    if not torch.cuda.is_available():
        h = C.c_void_p()
        assert L.tm_shard_comm_create(0, 1, 1000, C.byref(h)) != 0 and not h.value
        assert L.tm_last_error()
    code = C.c_int32(7)
    assert L.tm_shard_comm_status(None, C.byref(code)) != 0


def test_header_is_plain_c_and_a_c_host_links_against_the_library(built_lib, tmp_path):
    """include/transmil_b200.h is the boundary for hosts in any language: it must compile as strict C99 (no C++-isms, no torch or
    CUDA headers), and a C program linked against libtransmil_b200.so reaches the host-only entry points (no device needed)."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    src = tmp_path / "host.c"
    src.write_text(r'''
#include <stdio.h>
#include <string.h>
#include "transmil_b200.h"
int main(void) {
  tm_plan p;
  if (tm_make_plan(50000, &p) != TM_OK) return 1;
  if (p.G != 224 || p.S != 50177 || p.Sp != 50432 || p.P != 255 || p.l != 197) return 2;     /* SURVEY 8, cfg 3 */
  tm_weights w;
  memset(&w, 0, sizeof w);
  w.fc1_kind = TM_FC1_TOKENS; w.in_features = 768; w.n_classes = 2;
  size_t bytes = 0;
  if (tm_workspace_bytes(1, 100, &w, 0, &bytes) == TM_OK) return 3;                          /* NULL parameters are refused */
  if (strlen(tm_last_error()) == 0) return 4;
  if (tm_attn3v_tc_fwd_static(&p, &p, &p, 19, 256, (float*)&p, 0, 0) != TM_ERR_BAD_SHAPE) return 5;   /* beyond the static array */
  printf("%s\n", tm_version());
  return 0;
}
''')
    exe = tmp_path / "host"
    libdir = os.path.dirname(built_lib)
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                    "-L", libdir, "-ltransmil_b200", "-Wl,-rpath," + libdir], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "sm_100a" in r.stdout
