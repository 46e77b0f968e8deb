"""World-size-2 tests on CPU (gloo): the host-side logic of the N>1 path.
  * event ranges tile the event list, and rank-ordered concatenation of per-rank oracle output equals
    the single-process output (the `omp ordered` file order);
  * the multi-rank Trainer algorithm (exact u32/u64 partial sums all-reduced, fp64 finish with boundary
    flags, all-gather of the flagged pixels' samples, sequential float32 recurrence in rank order) is
    restated in numpy over torch.distributed/gloo and must reproduce the oracle's sequential Welford
    bit for bit -- the same algorithm csrc/nccl_glue.cu runs over NCCL."""
import os
import socket
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r"""
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
import oracle.cv2_oracle as o
from autobub3hs_b200 import shard, synth
out = sys.argv[2]
dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{sys.argv[3]}", rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, world = dist.get_rank(), dist.get_world_size()
run = synth.make_run(5, ["normal", "notrigger", "multi", "short5", "normal"], rows=66, cols=106, seed=4)
lo, hi = shard.event_range(len(run.events), rank, world)

# ---- Trainer: exact integer partials + flags + gathered fix-up -------------------------------------------
F32, F64 = np.float32, np.float64
planes = []
for cam in range(4):
    kept = []
    for ev in run.events[lo:hi]:
        fr = ev[cam]
        if len(fr) >= 2 and float(o.training_gate_entropy(fr[0], fr[1])) <= 0.0005:
            kept += [fr[0], fr[1]]
    st = np.stack(kept).astype(np.int64) if kept else np.zeros((0, 66, 106), np.int64)
    s = torch.from_numpy(st.sum(0)); q = torch.from_numpy((st * st).sum(0)); n = torch.tensor([len(kept)])
    for t in (s, q, n):
        dist.all_reduce(t)                       # integer sums: associative, so the result is rank-independent
    N = int(n.item()); s = s.numpy().astype(F64); q = q.numpy().astype(F64)
    mean = s / N
    M2 = np.maximum(q - s * s / N, 0.0) if N > 1 else np.zeros_like(s)
    var = M2 / max(N - 1, 1)
    sg = np.sqrt(var)
    # the flag rule of k4_finish_partials_kernel (csrc/k34_trainer.cu): worst-case float32 error of the recurrence
    u = 1.01 * 2.0 ** -24
    e_mean = 256.0 * u * (0.5 * N + 2.5)
    bv = (u * M2 * (N + 3.0) + 2.0 * e_mean * np.sqrt(2.0 * N * M2) + N * e_mean * e_mean) / max(N - 1, 1) + 1e-9
    rb, tail = np.sqrt(bv), 3.0 * u * sg + 1e-12
    with np.errstate(all="ignore"):
        e_up = np.where(sg > 0, np.minimum(bv / (2.0 * sg), rb), rb) + tail
        e_dn = np.where(sg > 0, np.minimum(bv / sg, rb), rb) + tail
    fm, fs = mean - np.floor(mean), sg - np.floor(sg)
    flag = ((M2 > 0) & ((fm < e_mean) | (1 - fm < e_mean) | (fs < e_dn) | (1 - fs < e_up))) | (N < 2)
    mean8 = mean.astype(np.uint8); sig8 = sg.astype(np.int64).astype(np.uint8)
    idx = np.flatnonzero(flag.ravel())
    counts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([len(kept)]))
    counts = [int(c.item()) for c in counts]
    stride = max(max(counts), 1)
    loc = np.zeros((len(idx), stride), np.uint8)
    if kept:
        loc[:, :len(kept)] = st.reshape(len(kept), -1)[:, idx].T
    allb = [torch.zeros((len(idx), stride), dtype=torch.uint8) for _ in range(world)]
    dist.all_gather(allb, torch.from_numpy(loc))
    for i, p in enumerate(idx):
        samples = np.concatenate([allb[r][i, :counts[r]].numpy() for r in range(world)])
        m, s8 = o.welford_mean_sigma([np.array([[v]], np.uint8) for v in samples])
        mean8.ravel()[p] = m[0, 0]; sig8.ravel()[p] = s8[0, 0]
    planes.append((mean8, sig8, N, int(flag.sum())))

# ---- detection on this rank's range with the replicated planes ------------------------------------------
full = o.train_all(run.events, 4)
ok = all(np.array_equal(p[0], t.avg) and np.array_equal(p[1], t.sigma) and p[2] == t.training_set_size for p, t in zip(planes, full))
text, _ = o.run_detect(run.run_id, run.event_numbers[lo:hi], run.events[lo:hi], full, run.cam_masks, run.bellows_masks)
parts = [None] * world
dist.all_gather_object(parts, text)
if rank == 0:
    ref, _ = o.run_detect(run.run_id, run.event_numbers, run.events, full, run.cam_masks, run.bellows_masks)
    merged = shard.merge_rank_outputs(o.write_header(), parts)
    open(out, "w").write(f"{int(ok)} {int(merged == o.write_header() + ref)} {sum(p[3] for p in planes)}")
else:
    open(out + ".1", "w").write(str(int(ok)))
dist.destroy_process_group()
"""


def test_event_ranges_tile():
    from autobub3hs_b200 import shard
    for n in (0, 1, 7, 20, 1000, 10000):
        for w in (1, 2, 3, 4, 8):
            rs = shard.all_ranges(n, w)
            assert rs[0][0] == 0 and rs[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(rs, rs[1:]))
            sizes = [b - a for a, b in rs]
            assert max(sizes) - min(sizes) <= 1


def test_world2_gloo_trainer_merge_and_ordered_output(tmp_path):
    w = str(tmp_path / "worker.py")
    open(w, "w").write(WORKER)
    out = str(tmp_path / "res.txt")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, w, ROOT, out, str(port)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    logs = [p.communicate(timeout=600)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), "\n".join(l[-3000:] for l in logs)
    planes_ok, text_ok, nflag = open(out).read().split()
    assert planes_ok == "1", "merged Trainer planes differ from the sequential float32 Welford"
    assert open(out + ".1").read() == "1"
    assert text_ok == "1", "rank-ordered concatenation differs from the single-process output"
    assert int(nflag) > 0, "the boundary fix-up was not exercised"
