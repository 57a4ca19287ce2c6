"""Multi-GPU correctness check of the partitioned solver over NCCL, one GPU per rank.  Launched by
tests/test_gpu_multi_nccl.py (pytest, skipped on a single-GPU box) or by hand under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29544 tests/nccl_partitioned_check.py

Every rank holds a slab of one 200,000-cell graded mesh (production tiles, 4096 cells); the
slabs' results are gathered and compared with (i) the same solver on a single rank and (ii) the
oracle's scipy loop (the reference's algorithm).  Prints one JSON line on rank 0; exits non-zero
on failure.  (The single-GPU emulation of this test, ranks over gloo sharing one GPU, is
tests/test_gpu_partitioned.py::test_cells_split_over_ranks.)"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fvm_b200 import PartitionedAdvectionDiffusionModel, workloads as W  # noqa: E402
from oracle import fvm_oracle as O  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    J, nsteps = 200000, 12
    f = W.graded_faces(J, 100.0)
    c = 0.5 * (f[1:] + f[:-1])
    h = np.diff(f)
    a = np.array([1.0, -0.6])
    d = np.abs(a) * h.max() / 2.0
    k = np.array([4.0, 9.0]) * h.min()
    w0 = 0.1 + np.stack([W.pulse(c), W.pulse(1.0 - c)])
    bc = dict(left_flux=0.2, right_flux=-0.1)
    m = PartitionedAdvectionDiffusionModel(f, a, d, k, "exponential", P=2, group=dist.group.WORLD)
    m.set_boundary_conditions(**bc)
    mine = m.run(w0, 0.5, nsteps)                                    # (2, j1-j0) on this rank's GPU
    pad = torch.zeros((2, m.n_local), dtype=torch.float64, device=mine.device)
    pad[:, :mine.shape[1]] = mine
    allp = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(allp, pad)
    ok, report = True, {}
    if rank == 0:
        got = torch.cat(allp, dim=1)[:, :J].cpu().numpy()
        single = PartitionedAdvectionDiffusionModel(f, a, d, k, "exponential", P=2)
        single.set_boundary_conditions(**bc)
        ws = single.run(w0, 0.5, nsteps).cpu().numpy()
        report["max_abs_vs_single_rank"] = float(np.max(np.abs(got - ws)))
        errs = []
        for p in range(2):
            asm = O.assemble(f, a[p], d[p], "exponential", **bc)
            want = O.theta_run_scipy(asm, k[p], 0.5, w0[p], nsteps, every=nsteps)[0]
            errs.append(float(np.max(np.abs(got[p] - want)) / np.max(np.abs(want))))
        report["max_rel_vs_oracle"] = max(errs)
        report.update(world=world, J=J, tiles_per_rank=m.nt, nsteps=nsteps, backend="nccl")
        ok = report["max_abs_vs_single_rank"] < 1e-13 and report["max_rel_vs_oracle"] < 1e-10
        report["ok"] = ok
        print(json.dumps(report), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
