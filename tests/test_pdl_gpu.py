"""GPU suite: programmatic dependent launch must not change a single bit.  The same bags run in three fresh processes --
TM_PDL=0 (ordinary launches), TM_PDL_X=0 (every kernel launched programmatically) and the default mask -- and the logits
and a checksum of every kept stage buffer have to be identical (a kernel that touched its inputs before
griddepcontrol.wait, or a lost dependency, shows up as a difference here)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import json, sys, torch
sys.path.insert(0, %r); sys.path.insert(0, %r)
from common import build_model, make_input
import transmil_b200 as T
out = {}
for D, n, B in ((1024, 20000, 1), (2048, 1000, 4), (1024, 6000, 1)):
    m = build_model(2, D).cuda()
    m.flags = T.capi.TM_FLAG_KEEP_STAGES
    x = make_input("rand", B, n, D, 77).cuda()
    with torch.no_grad():
        for rep in range(3):                       # back-to-back forwards: the next forward's first kernel follows the head kernel
            logits = m(x)
    torch.cuda.synchronize()
    rec = {"logits": logits.cpu().view(torch.int32).flatten().tolist()}
    for name in ("h0", "h1", "h2", "h3"):
        t = m.stage(name)
        rec[name] = [float(t.double().sum()), float(t.double().abs().sum())]
    out["%%d-%%d-%%d" %% (D, n, B)] = rec
print("RESULT " + json.dumps(out))
""" % (ROOT, os.path.join(ROOT, "tests"))


def _run(env_extra):
    env = dict(os.environ)
    env.pop("TM_PDL", None); env.pop("TM_PDL_X", None)
    env.update(env_extra)
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")][-1]
    return json.loads(line[len("RESULT "):])


def test_programmatic_launch_is_bit_identical_to_ordinary_launches(built_lib):
    plain = _run({"TM_PDL": "0"})
    every = _run({"TM_PDL_X": "0"})
    default = _run({})
    assert plain == every
    assert plain == default
