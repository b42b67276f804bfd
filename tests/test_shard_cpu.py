"""CPU tests of the N>1 (column-sharded) path: the sharding plan of the host flattener, and — with two `gloo`
ranks — the exchange + merge-in-shard-order logic that ggq_get_rows_sharded_device runs over NCCL on GPUs
(all-gather of partial offsets, all-gather-v of partial column ids, per-query concatenation in rank order).
The per-shard partial rows are produced here by the C oracle filtered to the shard's columns (no GPU), so what
is tested is exactly the host-side contract: shards own contiguous DFS runs, hence rank-order concatenation
equals BRWT::get_row's order (BRWT.cpp:45-50)."""
import glob
import os
import sys

import numpy as np
import pytest

from conftest import ROOT

GOLD = os.path.join(ROOT, "tests", "golden")
FIXTURES = sorted(glob.glob(os.path.join(GOLD, "brwt_*.npz")))


@pytest.mark.parametrize("path", FIXTURES)
def test_shard_plan_properties(ggq, path):
    g = np.load(path)
    data = g["stream"].tobytes()
    info = ggq.inspect_brwt(data)
    m = info["num_columns"]
    dfs = info["dfs_columns"]
    for R in (1, 2, 3, 4, 8, 16):
        owner, cut = ggq.shard_plan(data, R)
        assert owner.size == m
        if m == 0:
            continue
        assert owner.max() < R
        run = owner[dfs.astype(np.int64)]            # owners along the DFS (= get_row) order
        assert np.all(np.diff(run.astype(np.int64)) >= 0), "shards must own contiguous DFS runs"
        if R == 1:
            assert cut == 0 and np.all(owner == 0)
        else:
            assert 1 <= cut <= max(1, info["max_depth"])
            counts = np.bincount(owner, minlength=R)
            if m >= 4 * R and info["max_depth"] >= 2:
                assert counts.max() <= 2 * int(np.ceil(m / R)) + 1   # balanced by column count
        # the plan depends on the tree only: same answer again
        assert np.array_equal(ggq.shard_plan(data, R)[0], owner)


def _rank_main(rank, world, port, path, q):
    try:
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        sys.path.insert(0, ROOT)
        import torch
        import torch.distributed as dist
        import genome_graph_annotation_b200 as ggq
        dist.init_process_group("gloo", rank=rank, world_size=world)
        g = np.load(path)
        data = g["stream"].tobytes()
        n = int(g["stats"][0])
        eoff, ecols = g["row_offsets"].astype(np.int64), g["row_cols"]
        rows = np.random.default_rng(9).integers(0, max(n, 1), 777) if n else np.zeros(0, np.int64)
        owner, _ = ggq.shard_plan(data, world)
        # this rank's partial CSR: reference rows restricted to the owned columns, order kept
        p_off = np.zeros(rows.size + 1, np.int64)
        p_cols = []
        for t, r in enumerate(rows):
            row = ecols[eoff[r]:eoff[r + 1]]
            mine = row[owner[row.astype(np.int64)] == rank]
            p_cols.append(mine)
            p_off[t + 1] = p_off[t] + mine.size
        p_cols = np.concatenate(p_cols) if p_cols else np.zeros(0, np.uint64)
        # exchange: totals, offsets (all-gather), column ids (all-gather-v via padding to the max)
        tot = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(tot, torch.tensor([p_cols.size]))
        tot = [int(x) for x in tot]
        offs = [torch.zeros(rows.size + 1, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(offs, torch.from_numpy(p_off))
        pad = max(max(tot), 1)
        send = np.zeros(pad, np.int64)
        send[:p_cols.size] = p_cols.astype(np.int64)
        cols = [torch.zeros(pad, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(cols, torch.from_numpy(send))
        # merge: per query, segments in rank order (restatement of merge_count/merge_copy kernels)
        out_off = np.zeros(rows.size + 1, np.int64)
        out = []
        for t in range(rows.size):
            for r in range(world):
                o = offs[r].numpy()
                out.append(cols[r].numpy()[o[t]:o[t + 1]])
            out_off[t + 1] = out_off[t] + sum(int(offs[r][t + 1] - offs[r][t]) for r in range(world))
        out = np.concatenate(out).astype(np.uint64) if out else np.zeros(0, np.uint64)
        exp = np.concatenate([ecols[eoff[r]:eoff[r + 1]] for r in rows]) if rows.size else np.zeros(0, np.uint64)
        exp_off = np.concatenate([[0], np.cumsum([eoff[r + 1] - eoff[r] for r in rows])]).astype(np.int64)
        ok = np.array_equal(out, exp) and np.array_equal(out_off, exp_off)
        # slice mode (ggq_get_rows_sharded_slice_device): rank k merges only queries [n k / R, n (k + 1) / R); the
        # slices of all ranks, concatenated in rank order, are the whole answer
        q0, q1 = rows.size * rank // world, rows.size * (rank + 1) // world
        mine = []
        for t in range(q0, q1):
            for r in range(world):
                o = offs[r].numpy()
                mine.append(cols[r].numpy()[o[t]:o[t + 1]])
        mine = np.concatenate(mine).astype(np.int64) if mine else np.zeros(0, np.int64)
        ok = ok and np.array_equal(mine.astype(np.uint64), exp[int(exp_off[q0]):int(exp_off[q1])])
        sizes = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(sizes, torch.tensor([mine.size]))
        ok = ok and sum(int(x) for x in sizes) == exp.size
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, bool(ok), ""))
    except Exception as e:  # pragma: no cover
        q.put((rank, False, repr(e)))


@pytest.mark.parametrize("name", ["brwt_400x40_greedy_relax7.npz", "brwt_300x65_arity4.npz", "brwt_19x19_checker_relaxed.npz"])
def test_two_gloo_ranks_exchange_and_merge(name):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 400)
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, os.path.join(GOLD, name), q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    for rank, ok, err in res:
        assert ok, (rank, err)
