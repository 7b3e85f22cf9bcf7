"""torchrun --nproc-per-node N tools/score_multi_gpu.py : config 5 with the score-matrix rows
sharded over N GPUs (NCCL): one broadcast of T, one small reduce of the trial scores, EER on
rank 0.  Checks the result against the single-GPU path on rank 0 and times the sharded call."""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, ".")
from oracle import scoring_oracle as so  # noqa: E402  (checker only)
from speaker_verification_b200 import scoring  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
m = n = 4874
d = 512
E, T = so.synth_embeddings(m, n, d, seed=5)
ti, tj, lab = so.synth_trials(m, n, 37720, seed=5)
b = scoring.row_partition(m, world)
r0, r1 = b[rank], b[rank + 1]
E_blk = torch.from_numpy(E[r0:r1].copy()).to(dev)
T_dev = torch.from_numpy(T.copy()).to(dev) if rank == 0 else torch.zeros((n, d), dtype=torch.float32, device=dev)
d_ti, d_tj, d_lab = (torch.from_numpy(x.astype(np.int32)).to(dev) for x in (ti, tj, lab))
for _ in range(3):
    res = scoring.score_trials_sharded(E_blk, r0, m, T_dev, d_ti, d_tj, d_lab)
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
K = 10
e0.record()
for _ in range(K):
    res = scoring.score_trials_sharded(E_blk, r0, m, T_dev, d_ti, d_tj, d_lab)
e1.record()
torch.cuda.synchronize(); dist.barrier()
t = torch.tensor([e0.elapsed_time(e1) / K], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    eer, auc, full = res
    S = scoring.cosine_scores_dev(torch.from_numpy(E).to(dev), T_dev)
    single = S[d_ti.long(), d_tj.long()]
    eer1, auc1 = scoring.eer_auc_dev(single.contiguous(), d_lab)
    ref = so.cosine_scores(E[:64], T)
    print(f"world {world}: sharded trial scoring + EER {t.item():.3f} ms/call (max over ranks, device events); "
          f"scores equal single-GPU: {bool(torch.equal(full, single))}; eer {eer:.6f} (single {eer1:.6f}) "
          f"auc {auc:.6f} (single {auc1:.6f}); max |S - oracle| on 64 rows "
          f"{np.abs(S[:64].cpu().numpy() - ref).max():.2e}", flush=True)
dist.destroy_process_group()
