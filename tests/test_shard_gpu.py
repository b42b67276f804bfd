"""GPU parity of column sharding on ONE device: every shard of a matrix is loaded on
cuda:0, answers the batch with its owned columns only, and the merge kernel (ggq_merge_shards_device — the step
that follows the NCCL exchange in ggq_get_rows_sharded_device) must reproduce BRWT::get_row's answer, order
included, from the committed golden fixtures (reference streams + reference answers).  The real 2-rank NCCL
path is exercised by test_nccl_two_ranks when the box has two GPUs (scripts/shard_nccl_check.py)."""
import glob
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu
GOLD = os.path.join(ROOT, "tests", "golden")


def d2h(ptr, nbytes, dtype):
    from cuda.bindings import runtime as rt
    out = np.empty(nbytes // np.dtype(dtype).itemsize, dtype=dtype)
    if nbytes:
        (err,) = rt.cudaMemcpy(out.ctypes.data, ptr, nbytes, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)
        assert int(err) == 0, err
    return out


def h2d_rows(rows):
    import torch
    return torch.from_numpy(rows.view(np.int64)).cuda()


def filtered(off, cols, owned):
    """Reference CSR restricted to the owned columns, order kept."""
    nrows = off.size - 1
    keep = owned[cols.astype(np.int64)] if cols.size else np.zeros(0, bool)
    row_of = np.repeat(np.arange(nrows), np.diff(off.astype(np.int64)))
    cnt = np.bincount(row_of[keep], minlength=nrows) if nrows else np.zeros(0, np.int64)
    noff = np.zeros(off.size, np.uint64)
    noff[1:] = np.cumsum(cnt)
    return noff, cols[keep]


@pytest.mark.parametrize("n_shards", [2, 3, 4, 8])
@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "brwt_*.npz"))))
def test_sharded_rows_merge_to_reference(ggq, path, n_shards):
    g = np.load(path)
    data = g["stream"].tobytes()
    n, m = int(g["stats"][0]), int(g["stats"][1])
    eoff, ecols = g["row_offsets"], g["row_cols"]
    rows = np.arange(n, dtype=np.uint64)
    owner, cut = ggq.shard_plan(data, n_shards)
    shards = [ggq.GpuMatrix.load_brwt_shard(data, s, n_shards) for s in range(n_shards)]
    wss = [sm.workspace() for sm in shards]
    try:
        seen = np.zeros(m, np.int64)
        ptr_off, ptr_cols = [], []
        d_rows = h2d_rows(rows) if n else None
        for s, sm in enumerate(shards):
            sh, ns, owned = sm.shard_info()
            assert (sh, ns) == (s, n_shards)
            assert np.array_equal(owned, owner == s)
            seen += owned
            off, cols = sm.get_rows_csr(rows)
            foff, fcols = filtered(eoff, ecols, owned)
            assert np.array_equal(off, foff) and np.array_equal(cols, fcols), (s, n_shards)
            if n:
                o, c, t = wss[s].get_rows_device(d_rows.data_ptr(), n, 4)
                assert t == fcols.size
                ptr_off.append(o); ptr_cols.append(c)
        assert np.all(seen == 1)      # every column has exactly one owner
        if n:
            for col_bytes, dt in ((8, np.uint64), (4, np.uint32)):
                mo, mc, total = wss[0].merge_shards_device(n, ptr_off, ptr_cols, col_bytes)
                assert total == ecols.size
                assert np.array_equal(d2h(mo, (n + 1) * 8, np.uint64), eoff)
                assert np.array_equal(d2h(mc, total * col_bytes, dt).astype(np.uint64), ecols)
        # cells / columns live on the owner only
        coff, crows = g["col_offsets"], g["col_rows"]
        for c in range(0, m, max(1, m // 7)):
            for s, sm in enumerate(shards):
                if owner[c] == s:
                    assert np.array_equal(sm.get_column(c), crows[coff[c]:coff[c + 1]])
                    if n:
                        exp = np.isin(rows, crows[coff[c]:coff[c + 1]])
                        assert np.array_equal(sm.has_column(rows, c).astype(bool), exp)
                else:
                    with pytest.raises(ggq.GgqError):
                        sm.get_column(c)
    finally:
        for w in wss:
            w.close()
        for sm in shards:
            sm.close()


def test_synth_shards_equal_unsharded(ggq):
    """Device-built shards (harness of config C5) against the device-built whole tree, random batch with
    duplicates: the merged CSR is identical; images are smaller than the whole."""
    n, m, d, R = 40_000, 317, 0.038, 4
    whole = ggq.GpuMatrix.synthesize_brwt(n, m, d, seed=42, arity=4)
    shards = [ggq.GpuMatrix.synthesize_brwt_shard(n, m, d, s, R, seed=42, arity=4) for s in range(R)]
    wss = [sm.workspace() for sm in shards]
    try:
        rows = np.random.default_rng(5).integers(0, n, 30_000).astype(np.uint64)
        eoff, ecols = whole.get_rows_csr(rows)
        d_rows = h2d_rows(rows)
        po, pc = [], []
        for s, sm in enumerate(shards):
            assert sm.device_bytes < whole.device_bytes
            o, c, t = wss[s].get_rows_device(d_rows.data_ptr(), rows.size, 4)
            po.append(o); pc.append(c)
        assert sum(sm.num_relations for sm in shards) == whole.num_relations
        mo, mc, total = wss[0].merge_shards_device(rows.size, po, pc, 8)
        assert total == ecols.size
        assert np.array_equal(d2h(mo, (rows.size + 1) * 8, np.uint64), eoff)
        assert np.array_equal(d2h(mc, total * 8, np.uint64), ecols)
    finally:
        for w in wss:
            w.close()
        for sm in shards + [whole]:
            sm.close()


def test_nccl_two_ranks(ggq):
    """The collective path (ncclAllGather + grouped ncclBroadcast + merge) with one process per GPU."""
    if ggq.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29617", os.path.join(ROOT, "scripts", "shard_nccl_check.py")]
    p = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert p.returncode == 0 and "SHARD_NCCL_OK" in p.stdout, p.stdout[-4000:]
