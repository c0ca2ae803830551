"""8(e): host logic of the coset-sharded commit, exercised on CPU with a world_size-2 gloo group.

No GPU here: each rank takes ITS slice of the oracle's cap (standing in for the shard its GPU would compute) and the
real sharding.gather_cap() all-gathers the slices over gloo; the result must be the oracle's full cap on every rank.
The GPU side of the same path (shards computed by libnlc_commit.so) is tests/test_gpu_sharding.py.
"""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402


def test_plan_limits():
    S = graft.load_package().sharding
    assert S.check_plan(3, 4, 8) == 3 and S.check_plan(1, 4, 2) == 1 and S.check_plan(0, 0, 1) == 0
    with pytest.raises(S.ShardPlanError):
        S.check_plan(1, 4, 4)  # rate_bits = 1 shards two ways at most
    with pytest.raises(S.ShardPlanError):
        S.check_plan(3, 1, 4)  # fewer cap subtrees than shards
    with pytest.raises(S.ShardPlanError):
        S.check_plan(3, 4, 3)


def test_cosets_partition_the_domain():
    S = graft.load_package().sharding
    for r in range(0, 4):
        for G in [1 << s for s in range(r + 1)]:
            seen = sorted(k for g in range(G) for k in S.cosets_of(r, g, G))
            assert seen == list(range(1 << r))
            for g in range(G):
                lo, hi = S.leaf_range(5, r, g, G)
                n = 1 << 5
                # coset k owns leaves [bitrev_r(k) * n, +n): all of them must fall in the shard's range
                for k in S.cosets_of(r, g, G):
                    kr = int(format(k, f"0{r}b")[::-1], 2) if r else 0
                    assert lo <= kr * n and (kr + 1) * n <= hi


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        S = graft.load_package().sharding
        O = graft.load_oracle()
        rng = np.random.default_rng(7)
        W, log_n, r, cap = 5, 6, 2, 3
        values = O.random_field(rng, (W, 1 << log_n))
        ref = O.commit_from_values(values, r, cap)
        g, G = S.rank_and_world()
        S.check_plan(r, cap, G)
        local = ref.cap[S.cap_slice(cap, g, G)]
        full = S.gather_cap(local, cap)
        lo, hi = S.leaf_range(log_n, r, g, G)
        q.put((rank, bool(np.array_equal(full, ref.cap)), lo, hi))
    finally:
        dist.destroy_process_group()


def test_gather_cap_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [x[1] for x in res] == [True, True]
    assert (res[0][2], res[0][3], res[1][2], res[1][3]) == (0, 128, 128, 256)


def test_column_split_and_row_plan():
    S = graft.load_package().sharding
    assert S.column_split(10, 4) == [(0, 3), (3, 3), (6, 2), (8, 2)]
    assert S.column_split(8, 8) == [(j, 1) for j in range(8)]
    assert S.column_split(3, 4) == [(0, 1), (1, 1), (2, 1), (3, 0)]
    assert S.check_row_plan(4, 8) == 3  # rate_bits does not matter: starky's rate_bits 1 on 8 GPUs
    with pytest.raises(S.ShardPlanError):
        S.check_row_plan(2, 8)
    with pytest.raises(S.ShardPlanError):
        S.check_row_plan(4, 6)


def test_uniform_column_split():
    """The gathered coefficient matrix of the broadcast coset-sharded commit: equal strides, padding at the end."""
    S = graft.load_package().sharding
    assert S.uniform_column_split(135, 8) == (17, [(17 * p, 17) for p in range(7)] + [(119, 16)])
    assert S.uniform_column_split(3, 4) == (1, [(0, 1), (1, 1), (2, 1), (3, 0)])
    assert S.uniform_column_split(9, 8) == (2, [(0, 2), (2, 2), (4, 2), (6, 2), (8, 1), (9, 0), (9, 0), (9, 0)])
    for W in range(1, 40):
        for G in (1, 2, 4, 8):
            per, split = S.uniform_column_split(W, G)
            assert sum(nc for _, nc in split) == W and per * G >= W
            assert all(c0 + nc <= W and (nc == 0 or c0 == p * per) for p, (c0, nc) in enumerate(split))


def _row_worker(rank, world, port, q):
    """8e.3 host logic over gloo: the oracle plays the kernels; the exchange (pack_rows + all_to_all_single with
    ragged column splits) must hand every rank exactly its leaf rows, and the gathered cap is the commit's."""
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        S = graft.load_package().sharding
        O = graft.load_oracle()
        rng = np.random.default_rng(11)
        W, log_n, r, cap = 7, 6, 1, 3  # 7 columns over 2 ranks: 4 + 3
        values = O.random_field(rng, (W, 1 << log_n))
        ref = O.commit_from_values(values, r, cap)
        g, G = S.rank_and_world()
        S.check_row_plan(cap, G)
        c0, nc = S.column_split(W, G)[g]
        lde_cols = np.ascontiguousarray(ref.leaves[:, c0:c0 + nc].T)  # what nlc_lde_columns_dev returns
        send = torch.from_numpy(S.pack_rows(lde_cols, G).view(np.int64))
        rows = S.exchange_rows(send, W).numpy().view(np.uint64)
        lo, hi = S.leaf_range(log_n, r, g, G)
        ok_rows = bool(np.array_equal(rows, ref.leaves[lo:hi].T))
        # the tree over the shard (oracle in place of nlc_commit_from_lde_rows_dev), then the all-gather
        _, local_cap = O.merkle_new(np.ascontiguousarray(rows.T), cap - (G.bit_length() - 1))
        full = S.gather_cap(local_cap, cap)
        q.put((rank, ok_rows, bool(np.array_equal(full, ref.cap))))
    finally:
        dist.destroy_process_group()


def test_row_exchange_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_row_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True, True), (1, True, True)]
