"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): one process per GPU under torch.distributed.run,
NCCL over NVLink.  Both ways of sharding ONE commit must give the oracle's cap on every rank, and every shard's leaves
and digests must be the oracle's rows of that shard:
  8e.2  coset shards          (PolynomialBatch.from_values_sharded,     world <= 2^rate_bits)
  8e.3  column -> row shards  (PolynomialBatch.from_values_row_sharded, one all-to-all)
"""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np
sys.path.insert(0, os.environ["NLC_ROOT"])
import torch, torch.distributed as dist
import __graft_entry__ as graft
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
pkg, O = graft.load_package(), graft.load_oracle()
P2, S = pkg.plonky2, pkg.sharding
ctx = P2.Context(local)
rng = np.random.default_rng(5)
for W, log_n, r, cap in [(9, 8, 1, 4), (135, 10, 3, 4), (418, 9, 1, 4), (world, 6, 1, 3)]:
    values = O.random_field(rng, (W, 1 << log_n))
    ref = O.commit_from_values(values, r, cap)
    lo, hi = S.leaf_range(log_n, r, rank, world)
    dlo, dhi = (ref.digests.shape[0] // world) * rank, (ref.digests.shape[0] // world) * (rank + 1)
    for exchange in ("nccl", "p2p"):  # all-to-all of the scattered send buffer / NTT stores into the peers' buffers
        b = P2.PolynomialBatch.from_values_row_sharded(list(values), r, False, cap, ctx=ctx, exchange=exchange)
        assert np.array_equal(b.merkle_tree.cap, ref.cap), ("rows cap", exchange, W, log_n, r)
        assert np.array_equal(b.merkle_tree.leaves, ref.leaves[lo:hi]), ("rows leaves", exchange, W, log_n, r)
        assert np.array_equal(b.merkle_tree.digests, ref.digests[dlo:dhi]), ("rows digests", exchange, W, log_n, r)
        c0, nc = b.column_range
        assert np.array_equal(b.column_coeffs.cpu().numpy().view(np.uint64), ref.coeffs[c0:c0 + nc])
        b.free()
        del b
        dist.barrier()
    if world <= 1 << r:
        c = P2.PolynomialBatch.from_values_sharded(list(values), r, False, cap, ctx=ctx)
        assert np.array_equal(c.merkle_tree.cap, ref.cap), ("coset cap", W, log_n, r)
        assert np.array_equal(c.merkle_tree.leaves, ref.leaves[lo:hi])
        assert np.array_equal(c.merkle_tree.digests, ref.digests[dlo:dhi])
        # the same split with column-sharded upload + iFFT and the coefficient all-gather done by the iFFT's stores
        c2 = P2.PolynomialBatch.from_values_coset_sharded(list(values), r, False, cap, ctx=ctx)
        assert np.array_equal(c2.merkle_tree.cap, ref.cap), ("coset bcast cap", W, log_n, r)
        assert np.array_equal(c2.merkle_tree.leaves, ref.leaves[lo:hi])
        assert np.array_equal(c2.merkle_tree.digests, ref.digests[dlo:dhi])
        assert np.array_equal(c2.coeff_matrix, ref.coeffs)
        c2.free()
        del c2
        dist.barrier()
dist.barrier()
if rank == 0:
    print("multi-gpu parity ok, world", world)
dist.destroy_process_group()
'''


def test_sharded_commits_across_gpus(tmp_path):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU on this box")
    world = 1 << (min(n, 8).bit_length() - 1)
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, NLC_ROOT=ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(29400 + os.getpid() % 500), str(script)],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "multi-gpu parity ok" in r.stdout
