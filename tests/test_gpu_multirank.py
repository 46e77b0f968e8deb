"""Two processes, two GPUs: the Trainer's exact integer partials are all-reduced over NCCL and the
flagged pixels repaired from an all-gather of their samples -- planes must be bit-identical to the
oracle's sequential float32 Welford over the frames in global push order."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import sys, numpy as np
sys.path.insert(0, sys.argv[4])
from autobub3hs_b200 import capi
rank, world, idf = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
d = np.load(sys.argv[5])
frames, lo, hi = d["frames"], int(d["cuts"][rank]), int(d["cuts"][rank + 1])
rows, cols = frames.shape[1:]
with capi.Context(rows, cols, 1, 4, 1, device=rank) as ctx:
    comm = ctx.nccl_comm_create(idf, rank, world)
    ctx.train_begin(0)
    kept = 0
    for i in range(lo, hi, 2):
        _, k = ctx.train_add_event(0, frames[i], frames[i + 1])
        kept += 2 * int(k)
    ctx.train_partials(0)
    ctx.train_allreduce(0, comm)
    mean, sigma, n, st, nflag = ctx.train_finish_allreduced(0, comm)
    ctx.nccl_barrier(comm)
    np.savez(sys.argv[6] + f".{rank}.npz", mean=mean, sigma=sigma, n=n, st=st, nflag=nflag, kept=kept)
    ctx.nccl_comm_destroy(comm)
"""


def _ngpus():
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True).stdout
        return sum(1 for l in out.splitlines() if l.startswith("GPU "))
    except Exception:
        return 0


@pytest.mark.skipif(_ngpus() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("npairs,cut", [(7, 3), (2, 1), (30, 11)])
def test_trainer_two_ranks_exact(tmp_path, npairs, cut):
    import oracle.cv2_oracle as o
    rng = np.random.default_rng(npairs)
    rows, cols = 48, 80
    base = rng.integers(0, 250, (rows, cols)).astype(np.float32)
    frames = []
    for i in range(npairs):
        f0 = np.clip(base + rng.normal(0, 1.2, (rows, cols)), 0, 255).astype(np.uint8)
        f1 = f0.copy()                               # identical pair: passes the entropy gate
        frames += [f0, f1]
    frames = np.stack(frames)
    frames[:, :8] = 77                               # constant pixels: the recurrence is exact there, nothing to repair
    # every frame 10 but one pair at 10 + npairs: the mean is exactly 11 with a non-zero spread -> on an integer boundary, flagged
    frames[:, 8:16] = 10
    frames[:2, 8:16] = 10 + npairs
    data = str(tmp_path / "data.npz")
    np.savez(data, frames=frames, cuts=np.array([0, 2 * cut, 2 * npairs]))
    wfile = str(tmp_path / "worker.py")
    open(wfile, "w").write(WORKER)
    idf, out = str(tmp_path / "nccl.id"), str(tmp_path / "res")
    procs = [subprocess.Popen([sys.executable, wfile, str(r), "2", idf, ROOT, data, out], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    logs = [p.communicate(timeout=300)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(l[-2000:] for l in logs)
    mean_ref, sigma_ref = o.welford_mean_sigma(list(frames))
    for r in range(2):
        res = np.load(out + f".{r}.npz")
        assert int(res["n"]) == len(frames) and int(res["st"]) == 0
        assert int(res["nflag"]) >= 8 * cols
        assert np.array_equal(res["mean"], mean_ref), f"rank {r}"
        assert np.array_equal(res["sigma"], sigma_ref), f"rank {r}"
