"""MEASUREMENT TOOL: feature files -> logits.  Writes --slides synthetic bag files in the reference's layout
(`features [n, D]` chunked (1, D) + resizable, `coords [n, 2]`; tests/h5_writer.py, h5py is not in the image) under --dir, then
times the whole loop a user runs: `h5bag.iter_bags(paths, workers=W)` (file -> pinned host memory on reader threads) ->
`StreamedInference(model)` (H2D of the next bag under the current forward) -> logits on the host.  One JSON line per worker count.

    python tools/bench_files.py [--slides 24 --n 20000 --features 2048 --workers 0,2,4,8] [--f16]
"""
import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import h5_writer  # noqa: E402
import transmil_b200 as T  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--slides", type=int, default=24)
    ap.add_argument("--n", type=int, default=20000)
    ap.add_argument("--features", type=int, default=2048)
    ap.add_argument("--classes", type=int, default=2)
    ap.add_argument("--workers", default="0,2,4,8")
    ap.add_argument("--f16", action="store_true")
    ap.add_argument("--dir", default=None)
    a = ap.parse_args()
    d = a.dir or tempfile.mkdtemp(prefix="tm_bags_")
    rng = np.random.default_rng(0)
    paths = []
    for i in range(a.slides):
        n = int(a.n * rng.uniform(0.7, 1.3))
        x = rng.random((n, a.features), dtype=np.float32)
        x = x.astype(np.float16) if a.f16 else x
        c = rng.integers(0, 100000, (n, 2)).astype(np.int64)
        paths.append(h5_writer.write_h5(os.path.join(d, f"slide{i}.h5"),
                                        {"features": (x, dict(chunks=(1, a.features), maxshape=(None, a.features))), "coords": c}))
    nbytes = sum(os.path.getsize(p) for p in paths)
    torch.manual_seed(2021)
    model = T.TransMIL(n_classes=a.classes, in_features=a.features).eval().cuda()
    loop = T.StreamedInference(model)
    list(loop(T.h5bag.iter_bags(paths[:3])))                       # warm-up: library, workspace, pinned blocks, page cache
    for w in [int(v) for v in a.workers.split(",")]:
        best = None
        for _ in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            out = list(loop(T.h5bag.iter_bags(paths, workers=w)))
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        assert len(out) == len(paths) and all(torch.isfinite(o).all() for o in out)
        print(json.dumps({"tool": "files_to_logits", "slides": a.slides, "patches_about": a.n, "features": a.features,
                          "stored": "f16" if a.f16 else "f32", "reader_threads": w, "seconds": best, "slides_per_s": a.slides / best,
                          "file_GB_per_s": nbytes / best / 1e9, "host_cores": os.cpu_count(),
                          "timing": "host clock around the whole loop (page-cache-resident files), synchronize on both sides, best of 2",
                          "device": torch.cuda.get_device_name(0)}))


if __name__ == "__main__":
    main()
