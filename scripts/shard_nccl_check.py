"""Two (or more) ranks, one process per GPU: column-sharded get_rows through NCCL (ggq_get_rows_sharded_device)
must equal the unsharded answer.  Launched by tests/test_shard_gpu.py::test_nccl_two_ranks and by hand:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/shard_nccl_check.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import genome_graph_annotation_b200 as g  # noqa: E402


def d2h(ptr, nbytes, dtype):
    from cuda.bindings import runtime as rt
    out = np.empty(nbytes // np.dtype(dtype).itemsize, dtype=dtype)
    if nbytes:
        (err,) = rt.cudaMemcpy(out.ctypes.data, ptr, nbytes, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)
        assert int(err) == 0, err
    return out


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ids = [g.Comm.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    comm = g.Comm(ids[0], rank, world, local)
    n, m, d = 200_000, 500, 0.02
    sm = g.GpuMatrix.synthesize_brwt_shard(n, m, d, rank, world, seed=42, arity=4, device=local)
    ws = sm.workspace()
    gen = torch.Generator(device="cuda")
    gen.manual_seed(1234)   # the same batch on every rank
    for B in (1, 1000, 300_000):
        rows = torch.randint(0, n, (B,), device="cuda", dtype=torch.int64, generator=gen)
        off, cols, total = ws.get_rows_sharded_device(comm, rows.data_ptr(), B, 8)
        h_off, h_cols = d2h(off, (B + 1) * 8, np.uint64), d2h(cols, total * 8, np.uint64)
        if rank == 0:
            ref_m = g.GpuMatrix.synthesize_brwt(n, m, d, seed=42, arity=4, device=local)
            eoff, ecols = ref_m.get_rows_csr(rows.cpu().numpy().view(np.uint64))
            ref_m.close()
            assert np.array_equal(h_off, eoff) and np.array_equal(h_cols, ecols), B
        print("rank %d B=%d total=%d exchange %.3f ms %.1f MB" % (rank, B, total, ws.stat("exchange_ms"),
                                                                  ws.stat("exchange_bytes") / 1e6), flush=True)
    whole = g.GpuMatrix.synthesize_brwt(n, m, d, seed=42, arity=4, device=local)   # every rank checks its own slice

    def check_slice(tag, B):
        """ggq_get_rows_sharded_slice_device: this rank ends with queries [B*rank/world, B*(rank+1)/world) only."""
        rows = torch.randint(0, n, (B,), device="cuda", dtype=torch.int64, generator=gen)
        first, cnt, off, cols, total = ws.get_rows_sharded_slice_device(comm, rows.data_ptr(), B, 8)
        assert (first, cnt) == (B * rank // world, B * (rank + 1) // world - B * rank // world), (tag, first, cnt)
        h_off, h_cols = d2h(off, (cnt + 1) * 8, np.uint64), d2h(cols, total * 8, np.uint64)
        eoff, ecols = whole.get_rows_csr(rows[first:first + cnt].cpu().numpy().view(np.uint64))
        assert np.array_equal(h_off, eoff) and np.array_equal(h_cols, ecols), (tag, B)
        print("rank %d slice %s B=%d -> queries [%d, %d) ids=%d exchange %.3f ms %.1f MB" %
              (rank, tag, B, first, first + cnt, total, ws.stat("exchange_ms"), ws.stat("exchange_bytes") / 1e6), flush=True)

    for B in (0, 1, 7, 1000, 300_001):
        check_slice("nccl", B)
    # peer-memory exchange (ggq_comm_enable_peer): merge straight from the peers' buffers over NVLink; then a batch
    # beyond the pinned capacity, which must fall back to the NCCL exchange on every rank — same answers throughout
    comm.enable_peer(ws, 400_000, 400_000 * 8)
    assert comm.peer_active
    for B in (0, 1, 7, 1000, 300_001):
        check_slice("peer", B)
    assert comm.peer_active
    for B, expect_peer in ((1, True), (1000, True), (300_000, True), (300_000, True), (0, True), (500_000, False), (1000, False)):
        rows = torch.randint(0, n, (B,), device="cuda", dtype=torch.int64, generator=gen)
        off, cols, total = ws.get_rows_sharded_device(comm, rows.data_ptr(), B, 8)
        h_off, h_cols = d2h(off, (B + 1) * 8, np.uint64), d2h(cols, total * 8, np.uint64)
        assert comm.peer_active == expect_peer, (B, comm.peer_active)
        if rank == 0:
            eoff, ecols = whole.get_rows_csr(rows.cpu().numpy().view(np.uint64))
            assert np.array_equal(h_off, eoff) and np.array_equal(h_cols, ecols), ("peer", B)
        print("rank %d peer=%d B=%d total=%d exchange %.3f ms %.1f MB" % (rank, expect_peer, B, total, ws.stat("exchange_ms"),
                                                                          ws.stat("exchange_bytes") / 1e6), flush=True)
    # count_labels across GPUs: column shards count their own columns over the whole batch, replicas count their slice
    # of the batch; one all-reduce over num_columns counters gives the histogram of the batch on every rank
    B = 200_000
    rows = torch.randint(0, n, (B,), device="cuda", dtype=torch.int64, generator=gen)
    np_rows = rows.cpu().numpy().view(np.uint64)
    expect = None
    if rank == 0:
        expect = whole.count_labels(np_rows)
    cnt = d2h(ws.count_labels_allreduce_device(comm, rows.data_ptr(), B), m * 8, np.uint64)
    replica = g.GpuMatrix.synthesize_brwt(n, m, d, seed=42, arity=4, device=local)
    rws = replica.workspace()
    lo, hi = B * rank // world, B * (rank + 1) // world
    cnt2 = d2h(rws.count_labels_allreduce_device(comm, rows.data_ptr() + lo * 8, hi - lo), m * 8, np.uint64)
    if rank == 0:
        assert np.array_equal(cnt, expect), "sharded count_labels + all-reduce"
        assert np.array_equal(cnt2, expect), "replica count_labels + all-reduce"
        print("count_labels all-reduce ok (shards and replicas)", flush=True)
    rws.close(); replica.close()
    if whole is not None:
        whole.close()
    dist.barrier()
    ws.close(); sm.close(); comm.close()
    dist.destroy_process_group()
    if rank == 0:
        print("SHARD_NCCL_OK")


if __name__ == "__main__":
    main()
