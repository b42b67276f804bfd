#!/usr/bin/env python
"""Headline benchmark (BASELINE.json: "Multi-BRWT get_rows rows/s at 1/2/4/8 B200").

Workload (config 4 of BASELINE.json, SURVEY.md §8d "C4"): Kingsford-shaped synthetic matrix
3.7e9 x 2652, density 0.3 %, defined implicitly by the portable generator
M[i][j] = mix64(42*0x9E3779B97F4A7C15 + i*m + j) < floor(d*2^64); BRWT built on the device with the basic
partitioner, arity 4 (shape-equivalent of Multi-BRWT, measured avg arity ~4); one step = one batch of
2^24 uniformly random row ids per GPU answered by get_rows (CSR out).  Replicas, batch-sharded: every
GPU holds the whole structure, no collective on the query path => "scaling": "weak".

  python bench.py --gpus N --steps K --warmup W          # ours (libggq.so, CUDA)
  python bench.py --impl reference ...                   # the reference's own BRWT::get_row on host cores

`value`  : device-resident rows/s (row ids and CSR in HBM), CUDA events on the query stream, max over ranks.
`e2e`    : same metric through the host-buffer C-ABI call ggq_get_rows_into_u32 (pinned host row ids in, host
           CSR offsets + u32 column ids out; H2D and D2H inside the timed region); `e2e_u64`: the same through
           ggq_get_rows_into (u64 column ids, twice the PCIe bytes).
`roofline`: brwt_rows_kernel (the descent), algorithmic bytes per SURVEY.md §8(d) / its event-timed duration.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "multi_brwt_get_rows_rows_per_s"
UNIT = "rows/s"
SEED = 42
L2_BYTES = 126e6


_REAL_STDOUT = None


def quiet_stdout():
    """NCCL (and friends) write banners such as 'NCCL version ...' to fd 1; the driver wants exactly ONE JSON
    line there.  Route fd 1 to stderr for the whole run and keep the real stdout for emit()."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=float, default=3.7e9, help="matrix rows (C4: 3.7e9)")
    ap.add_argument("--cols", type=int, default=2652)
    ap.add_argument("--density", type=float, default=0.003)
    ap.add_argument("--arity", type=int, default=4)
    ap.add_argument("--batch", type=int, default=1 << 24, help="row ids per GPU per step")
    ap.add_argument("--cpu-rows", type=float, default=1e6, help="rows of the sub-shape the CPU reference is built on")
    ap.add_argument("--workload", default="c4", choices=["c4", "c5"],
                    help="c4: Kingsford-shaped, replicas (headline); c5: RefSeq-shaped 1e9 x 3173 d=3.8%%, column-sharded "
                         "over all ranks with the NCCL exchange (needs --gpus >= 2)")
    ap.add_argument("--tree", default="basic", choices=["basic", "greedy_relax7"],
                    help="basic: BRWTBottomUpBuilder with get_basic_partitioner(arity) (the headline, SURVEY.md 8d); "
                         "greedy_relax7: the reference's Multi-BRWT recipe (greedy partitioner + relax(7)) on the same matrix")
    ap.add_argument("--spot-rows", type=int, default=1 << 20, help="rows of the full-scale parity sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-bitvector", action="store_true", help="skip the rank1/select1/get block (C3)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--c5-rows", type=float, default=1e9)
    ap.add_argument("--c5-cols", type=int, default=3173)
    ap.add_argument("--c5-density", type=float, default=0.038)
    ap.add_argument("--c5-batch", type=int, default=1 << 22, help="row ids per step (the same batch on every rank)")
    ap.add_argument("--c5-spot-rows", type=int, default=1 << 14)
    ap.add_argument("--c5-full", action="store_true", help="c5: every rank merges the whole batch instead of its slice")
    ap.add_argument("--no-c5", action="store_true", help="multi-GPU default run: skip the column-sharded C5 block")
    ap.add_argument("--no-peer", action="store_true",
                    help="c5: exchange partial rows with NCCL instead of merging from peer memory over NVLink")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 20 ms.  The sampler is started before the warm-up steps
    (nvidia-smi needs up to a second to come up on an 8-GPU box) and only the samples stamped inside the timed
    region are kept; if the region was too short to catch any, the samples of warm-up + timed region (the same
    load) are reported and `window` says so."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        self.index = index
        self.t0 = self.t1 = None

    def mark_begin(self):
        import datetime
        self.t0 = datetime.datetime.now()

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "20"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        import datetime
        self.t1 = datetime.datetime.now()
        if self.p is not None:
            self.p.terminate()
            try:
                self.p.wait(timeout=5)
            except Exception:
                self.p.kill()
        self.f.flush()
        self.f.seek(0)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = []
        for line in self.f.read().splitlines():
            parts = [x.strip() for x in line.split(",")]
            if len(parts) < 8:
                continue
            try:
                ts = datetime.datetime.strptime(parts[0], "%Y/%m/%d %H:%M:%S.%f")
                rows.append((ts, float(parts[1]), float(parts[2]),
                             [nm for nm, v in zip(names, parts[4:8]) if v.lower().startswith("active")]))
            except ValueError:
                continue
        self.f.close()
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        window = "timed"
        inside = [r for r in rows if self.t0 is None or self.t0 <= r[0] <= self.t1]
        if not inside:
            inside, window = rows, "warmup+timed"
        if not inside:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        reasons = set()
        for r in inside:
            reasons.update(r[3])
        return {"sm_mhz": float(np.median([r[1] for r in inside])), "sm_max_mhz": float(max(r[2] for r in inside)),
                "reasons": sorted(reasons), "samples": len(inside), "window": window}


# --------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the reference's own BRWT::get_row (BRWT.cpp:26-53), unmodified, under
# `omp parallel for` over row ids (oracle/ref_capi.cpp) on all host threads, on an n'-row sub-shape of
# the same generator (the reference builder cannot make 3.7e9 rows: SURVEY.md §7b(d)).
# --------------------------------------------------------------------------------------------
def host_threads():
    """Host threads this process may run on.  NOT omp_get_max_threads(): torch.distributed.run exports
    OMP_NUM_THREADS=1 to every rank, which would time the reference on one core (round-1 verdict)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def build_cpu_reference(args):
    threads = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(threads)   # before libgomp initialises (it is loaded with libggref.so)
    from oracle import ref_binding as R  # checker / baseline only
    import genome_graph_annotation_b200 as g
    if not R.available():
        return None
    n = int(args.cpu_rows)
    thr = g.density_threshold(args.density)
    cw = R.gen_implicit(n, args.cols, thr, SEED)
    rm = R.RefMatrix.build_brwt(cw, n, arity=args.arity, threads=threads)
    del cw
    return R, rm, n, threads


def cpu_sample_rate(rm, n, threads, seconds):
    """rows/s of omp-parallel reference get_row on a bounded random sample sized for ~`seconds`."""
    rng = np.random.default_rng(SEED)
    probe = rng.integers(0, n, 1 << 14).astype(np.uint64)
    s, _ = rm.time_get_rows(probe, threads)
    rate = probe.size / max(s, 1e-9)
    cnt = int(min(1 << 22, max(1 << 14, rate * seconds)))
    rows = rng.integers(0, n, cnt).astype(np.uint64)
    s, _ = rm.time_get_rows(rows, threads)
    return cnt / s, cnt, s


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    built = build_cpu_reference(args)
    line = {"metric": METRIC, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic"}
    if built is None:
        line["unavailable"] = "oracle/_ref/libggref.so (the unmodified reference) is not built on this box"
        emit(line)
        return
    R, rm, n, threads = built
    rng = np.random.default_rng(SEED)
    probe = rng.integers(0, n, 1 << 14).astype(np.uint64)
    s, _ = rm.time_get_rows(probe, threads)
    rate = probe.size / max(s, 1e-9)
    budget = 90.0 / max(1, args.steps + args.warmup)
    cnt = int(min(1 << 22, max(1 << 12, rate * budget)))
    times = []
    for it in range(args.warmup + args.steps):
        rows = rng.integers(0, n, cnt).astype(np.uint64)
        s, _ = rm.time_get_rows(rows, threads)
        if it >= args.warmup:
            times.append(s)
    ms = 1e3 * float(np.mean(times))
    v = cnt / (ms / 1e3)
    sample = ("%d random row ids per step on a %d x %d d=%g sub-shape of the same generator (reference builder, "
              "basic partitioner arity %d); unmodified BRWT::get_row under omp parallel for" %
              (cnt, n, args.cols, args.density, args.arity))
    line.update({"value": v, "ms_per_step": ms,
                 "config": workload_config(args, args.gpus),
                 "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "reference", "sample": sample,
                                  "cpu_sub_shape_rows": n,
                                  "why_sub_shape": "the reference builder materialises every column on the host: 3.7e9 x 2652 "
                                                   "cannot be built; 1e6 rows (node vectors cache-resident, which favours the "
                                                   "CPU) keeps the arm within its time budget (SURVEY.md 8d allows 1e6)"},
                 "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    emit(line)


def workload_config(args, n_gpus):
    tree = ("arity-%d basic partitioner" % args.arity if args.tree == "basic"
            else "greedy partitioner + relax(7) (the reference's Multi-BRWT recipe, built on the device)")
    c = {"workload": "C4 Kingsford-shaped synthetic BRWT %.3g x %d density %g, %s; "
                     "get_rows over %d uniformly random row ids per GPU per step" %
                     (args.rows, args.cols, args.density, tree, args.batch),
         "rows": int(args.rows), "cols": args.cols, "density": args.density, "arity": args.arity,
         "batch_rows_per_gpu": args.batch, "seed": SEED,
         "parallelism": "replicas x%d, batch-sharded, no data-path collective" % n_gpus,
         "l2": "inputs larger than L2 (random gathers over the whole device image)"}
    return c


def bitvector_block(cpu):
    """rank1 / select1 / get on 2^32-bit bit_vector_{stat,sd,rrr} at d = 1 % and 50 % (scripts/bench_bv.py holds the
    full C3 sweep): G ops/s, algorithmic GB/s per SURVEY.md 8(d) against the measured HBM peak, and the unmodified
    reference's ops/s on all host threads (2^28-bit prefix of the same generator, 2^20 answers compared)."""
    from scripts import bench_bv
    t0 = time.time()
    peak, res = bench_bv.measure(["stat", "sd", "rrr"], [0.01, 0.5], log2_bits=32, log2_queries=26, log2_cpu_bits=28, cpu=cpu)
    # Random single-sector probes are bounded by the memory system's transaction rate, not by bytes: 43-45 G independent
    # probes/s over 64-120 GiB footprints (scripts/mb_gather.cu, profiles/r02/mb_gather_footprints_64_124GiB.txt).
    # `probe_frac` = G ops/s x dependent DRAM probes per op of this layout / that ceiling (samples live in L2).
    gather_ceiling = 44.0
    probes = {"stat": {"get": 1, "rank1": 1, "select1": 2}, "sd": {"get": 3, "rank1": 3, "select1": 3},
              "rrr": {"get": 2, "rank1": 2, "select1": 3}}
    out = {"n_bits": 1 << 32, "queries": 1 << 26, "peak_hbm_gbs": peak, "random_probe_ceiling_gprobes_per_s": gather_ceiling,
           "seconds": None, "results": []}
    for r in res:
        e = {"kind": r["kind"], "density": r["density"], "device_bytes": r["device_bytes"], "residency": r["residency"]}
        for op, o in r["ops"].items():
            e[op] = {"gops_per_s": o["ops_per_s"] / 1e9, "bytes_per_op": o["algorithmic_bytes_per_op"],
                     "roofline_frac": o["frac_of_measured_hbm_peak"], "dependent_probes_per_op": probes[r["kind"]][op],
                     "probe_frac": o["ops_per_s"] / 1e9 * probes[r["kind"]][op] / gather_ceiling}
            if "cpu_reference_ops_per_s" in o:
                e[op]["cpu_gops_per_s"] = o["cpu_reference_ops_per_s"] / 1e9
        if "parity_vs_reference_2^20_samples" in r:
            e["identical_to_reference"] = all(r["parity_vs_reference_2^20_samples"].values())
            e["cpu_threads"] = r["cpu"]["threads"]
        out["results"].append(e)
    out["seconds"] = time.time() - t0
    return out


def direct_rows(rows, m, density, threads=None):
    """CSR (counts, column ids) of the given rows by direct evaluation of the implicit matrix
    M[i][j] = mix64(seed*K + i*m + j) < floor(d*2^64) (SURVEY.md 8d portable generator) in numpy — independent of
    the engine and of the oracle.  Columns ascending per row, which for the basic partitioner IS the DFS order of
    BRWT::get_row (BRWT.cpp:45-50).  Chunks of rows are evaluated on `threads` host threads (numpy releases the GIL)."""
    import genome_graph_annotation_b200 as g
    from concurrent.futures import ThreadPoolExecutor
    thr = np.uint64(g.density_threshold(density))
    base = np.uint64((SEED * g.SEED_MUL) % 2**64)
    jj = np.arange(m, dtype=np.uint64)[None, :]
    step = 2048

    def chunk(lo):
        r = rows[lo:lo + step]
        with np.errstate(over="ignore"):
            hit = g.mix64(base + r[:, None] * np.uint64(m) + jj) < thr
        ri, ci = np.nonzero(hit)
        return np.bincount(ri, minlength=r.size).astype(np.uint64), ci.astype(np.uint64)

    with ThreadPoolExecutor(max_workers=threads or host_threads()) as ex:
        parts = list(ex.map(chunk, range(0, rows.size, step)))
    counts = np.concatenate([p[0] for p in parts]) if parts else np.zeros(0, np.uint64)
    cols = np.concatenate([p[1] for p in parts]) if parts else np.zeros(0, np.uint64)
    return counts, cols


def spot_check(gm, rows, m, density, what="get_rows_csr"):
    t0 = time.time()
    s_off, s_cols = gm.get_rows_csr(rows)
    counts, cols = direct_rows(rows, m, density)
    dfs = gm.dfs_columns().astype(np.int64)
    if not np.array_equal(dfs, np.arange(m)):
        # a tree whose DFS leaf order is not the column order (greedy partitioner): every row must list its columns by
        # increasing DFS rank — re-order the expectation that way
        rank = np.empty(m, np.int64)
        rank[dfs] = np.arange(m)
        row_of = np.repeat(np.arange(rows.size), counts.astype(np.int64))
        order = np.lexsort((rank[cols.astype(np.int64)], row_of))
        cols = cols[order]
    same = (np.array_equal(np.diff(s_off.astype(np.int64)), counts.astype(np.int64)) and
            np.array_equal(np.asarray(s_cols, dtype=np.uint64), cols))
    if not same:
        raise RuntimeError("bench.py: rows of the measured matrix differ from direct evaluation")
    return {"rows": int(rows.size), "column_ids": int(cols.size), "identical": True, "seconds": time.time() - t0,
            "against": "direct numpy evaluation of the implicit matrix (SURVEY.md 8d generator) on the full-size "
                       "matrix: per-row counts and ids in DFS order, every row of the sample"}


# --------------------------------------------------------------------------------------------
# ours
# --------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import genome_graph_annotation_b200 as g

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or g.device_count() < 1:
        raise RuntimeError("bench.py: no CUDA device; the engine has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    n = int(args.rows)
    B = args.batch
    K, W = args.steps, args.warmup
    t0 = time.time()
    if args.tree == "greedy_relax7":
        gm = g.GpuMatrix.synthesize_brwt_greedy(n, args.cols, args.density, seed=SEED, relax=7, device=local)
    else:
        gm = g.GpuMatrix.synthesize_brwt(n, args.cols, args.density, seed=SEED, arity=args.arity, device=local)
    torch.cuda.synchronize()
    build_s = time.time() - t0
    ws = gm.workspace()
    stream = torch.cuda.ExternalStream(ws.stream, device=torch.device("cuda", local))
    # nvidia-smi needs up to a second to come up: start it now, long before the timed region (the timed steps of the
    # coded descent take ~0.1-0.2 s in all)
    clocks = ClockSampler(local)
    clocks.start()

    # distinct batches of row ids (seeded per rank); cycled if K + W exceeds the pool
    n_pool = min(K + W, 6)
    gen = torch.Generator(device="cuda")
    batches = []
    for b in range(n_pool):
        gen.manual_seed(SEED + 1000 * rank + b)
        batches.append(torch.randint(0, n, (B,), device="cuda", dtype=torch.int64, generator=gen))
    torch.cuda.synchronize()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def device_step(i):
        return ws.get_rows_device(batches[i % n_pool].data_ptr(), B, 8)

    # ---- the reference algorithm's work on these batches: V_t = (node, row) touches of BRWT::get_row,
    # counted by the plain descent (leaf codes off; they skip most of the touches the reference performs).  Untimed.
    codes = gm.leaf_codes
    touch_of, plain_ms = {}, []
    if codes["in_use"]:
        gm.set_option("leaf_codes", 0)
        for b in range(n_pool):
            device_step(b)
            touch_of[b] = int(ws.stat("touches"))
            plain_ms.append(ws.stat("kernel_ms"))
        gm.set_option("leaf_codes", 1)

    # ---- device-resident number ----------------------------------------------------------
    for i in range(W):
        device_step(i)
    barrier()
    clocks.mark_begin()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kern_ms, touches, probes, ids, launches, code_probes, code_esc = [], 0, 0, 0, 0, 0, 0
    with torch.cuda.stream(stream):
        e0.record()
    for i in range(K):
        _, _, total = device_step(W + i)
        kern_ms.append(ws.stat("kernel_ms"))
        touches += touch_of.get((W + i) % n_pool, int(ws.stat("touches")))
        probes += int(ws.stat("probes"))
        code_probes += int(ws.stat("code_probes"))
        code_esc += int(ws.stat("code_escapes"))
        launches += int(ws.stat("launches"))
        ids += total
    with torch.cuda.stream(stream):
        e1.record()
    barrier()
    clk = clocks.stop()
    ms_total = e0.elapsed_time(e1)
    t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / K
    value = world * B / (ms_step / 1e3)

    # ---- full-scale spot check (outside every timed region; SURVEY.md 8d "C4 / C5 concrete input"): every row of a
    # 2^20-row sample of the measured batch against direct evaluation of the implicit matrix
    spot = spot_check(gm, batches[0][:args.spot_rows].cpu().numpy().view(np.uint64), args.cols, args.density)

    # ---- the other calls of the BinaryMatrix interface on the same matrix (host to host, outside the headline):
    # get_column (BRWT.cpp:55-85) for a few columns, get (BRWT.cpp:9-24) for 2^22 random cells
    extra = None
    if not args.no_e2e:
        pick = [7, args.cols // 3, args.cols - 1]
        h_col = torch.empty(max(gm.get_column_count(c) for c in pick) + 1, dtype=torch.int64).pin_memory()
        np_col = h_col.numpy().view(np.uint64)
        for c in pick:   # warm-up: the result buffer grows to the largest column once
            gm.get_column_into(c, np_col)
        tcol = 1e30
        for rep in range(3):
            tw = time.perf_counter()
            p_out = sum(int(gm.get_column_into(c, np_col)) for c in pick)
            tcol = min(tcol, (time.perf_counter() - tw) / len(pick))
        del h_col
        nq = 1 << 22
        cr = batches[1 % n_pool][:nq].cpu().numpy().view(np.uint64)
        cc = (cr * np.uint64(2654435761) % np.uint64(args.cols)).astype(np.uint64)
        gm.get_cells(cr, cc)   # warm-up at full size (scratch buffers grow once)
        tw = time.perf_counter()
        cells = gm.get_cells(cr, cc)
        tcell = time.perf_counter() - tw
        # one query sequence at a time, the way annograph classify calls without cross-query batching:
        # 300 consecutive-ish row ids (a read's k-mers) -> get_labels(ratio 0.25), and a scalar get_row
        q300 = np.ascontiguousarray(cr[:300])
        for _ in range(20):
            gm.get_labels(q300, 0.25); gm.get_row(int(cr[0]))
        tw = time.perf_counter()
        for i in range(200):
            gm.get_labels(cr[i * 300:(i + 1) * 300], 0.25)
        t_lab = (time.perf_counter() - tw) / 200
        tw = time.perf_counter()
        for i in range(200):
            gm.get_row(int(cr[i]))
        t_row = (time.perf_counter() - tw) / 200
        extra = {"small_calls": {"get_labels_300_indices_us": t_lab * 1e6, "get_row_us": t_row * 1e6,
                                 "note": "python ctypes call overhead included"},
                 "get_column": {"api": "ggq_get_column_into (pinned host buffer)", "ms_per_column": tcol * 1e3, "rows_out_per_column": p_out / len(pick),
                                "rows_out_per_s": p_out / len(pick) / tcol},
                 "get": {"cells": nq, "cells_per_s": nq / tcell, "set_fraction": float(np.mean(cells))}}

    # ---- end to end through the host-buffer C ABI ------------------------------------------
    e2e = e2e_u32 = e2e_u64 = None
    if not args.no_e2e:
        cap = int(ids / K * 1.05) + 1024
        h_rows = [b.cpu().pin_memory() for b in batches[:min(n_pool, 3)]]
        h_off = torch.empty(B + 1, dtype=torch.int64).pin_memory()
        np_rows = [r.numpy().view(np.uint64) for r in h_rows]
        np_off = h_off.numpy().view(np.uint64)
        Ke, We = max(3, min(K, 5)), 2

        def host_api(np_cols, id_bytes, api):
            for i in range(We):
                gm.get_rows_into(np_rows[i % len(np_rows)], np_off, np_cols)
            barrier()
            tw = time.perf_counter()
            d2h = 0
            for i in range(Ke):
                total = gm.get_rows_into(np_rows[(We + i) % len(np_rows)], np_off, np_cols)
                d2h += (B + 1) * 8 + total * id_bytes
            torch.cuda.synchronize()
            te = torch.tensor([time.perf_counter() - tw], device="cuda", dtype=torch.float64)
            if dist is not None:
                dist.all_reduce(te, op=dist.ReduceOp.MAX)
            return {"value": world * B * Ke / float(te.item()), "unit": UNIT, "h2d_bytes_per_step": B * 8,
                    "d2h_bytes_per_step": d2h // Ke, "steps": Ke, "api": api}

        # headline: the compact call — u32 row ids in (the matrix has < 2^32 rows), u8 per-row counts + u16 column ids
        # out (< 65536 columns): 22 instead of 56 bytes per row over the host link, same rows (tests/test_synth_gpu.py)
        compact_ok = n <= 2**32 and args.cols <= 65536
        if compact_ok:
            h_rows32 = [torch.from_numpy(r.astype(np.uint32).view(np.int32)).pin_memory() for r in np_rows]
            np_rows32 = [r.numpy().view(np.uint32) for r in h_rows32]
            h_cnt8 = torch.empty(B, dtype=torch.uint8).pin_memory()
            h_cols16 = torch.empty(cap, dtype=torch.int16).pin_memory()
            np_cnt8, np_cols16 = h_cnt8.numpy(), h_cols16.numpy().view(np.uint16)

            def compact_api():
                for i in range(We):
                    gm.get_rows_compact(np_rows32[i % len(np_rows32)], np_cnt8, np_cols16)
                barrier()
                tw = time.perf_counter()
                d2h = 0
                for i in range(Ke):
                    total = gm.get_rows_compact(np_rows32[(We + i) % len(np_rows32)], np_cnt8, np_cols16)
                    d2h += B + total * 2
                torch.cuda.synchronize()
                te = torch.tensor([time.perf_counter() - tw], device="cuda", dtype=torch.float64)
                if dist is not None:
                    dist.all_reduce(te, op=dist.ReduceOp.MAX)
                return {"value": world * B * Ke / float(te.item()), "unit": UNIT, "h2d_bytes_per_step": B * 4,
                        "d2h_bytes_per_step": d2h // Ke, "steps": Ke,
                        "api": "ggq_get_rows_compact (host u32 row ids in; host u8 per-row counts + u16 column ids out)"}

            e2e = compact_api()
            # the timed call's last answer against the u32-id call on the same batch (outside the timed region)
            last = (We + Ke - 1) % len(np_rows32)
            tot_c = gm.get_rows_compact(np_rows32[last], np_cnt8, np_cols16)
            chk_cols = np.empty(cap, np.uint32)
            tot_r = gm.get_rows_into(np_rows[last], np_off, chk_cols)
            if tot_c != tot_r or not np.array_equal(np_cnt8.astype(np.int64), np.diff(np_off.astype(np.int64))) or \
                    not np.array_equal(np_cols16[:tot_c], chk_cols[:tot_r].astype(np.uint16)):
                raise RuntimeError("bench.py: ggq_get_rows_compact differs from ggq_get_rows_into_u32")
            e2e["identical_to_get_rows_into_u32"] = True
            del h_cnt8, h_cols16, chk_cols
        h_cols32 = torch.empty(cap, dtype=torch.int32).pin_memory()
        e2e_u32 = host_api(h_cols32.numpy().view(np.uint32), 4,
                           "ggq_get_rows_into_u32 (host u64 row ids in, host CSR offsets + u32 column ids out)")
        if e2e is None:
            e2e = e2e_u32
        del h_cols32
        h_cols = torch.empty(cap, dtype=torch.int64).pin_memory()
        e2e_u64 = host_api(h_cols.numpy().view(np.uint64), 8,
                           "ggq_get_rows_into (host row ids in, host CSR offsets + u64 column ids out)")
        del h_cols

    # ---- the call `annograph classify` really makes: count_labels (annotate_static.cpp:149-162), fused on device ----
    cl = None
    if not args.no_e2e:
        h_rows_cl = batches[0].cpu().pin_memory().numpy().view(np.uint64)
        gm.count_labels(h_rows_cl)
        barrier()
        tw = time.perf_counter()
        Kc = 3
        for i in range(Kc):
            cnt = gm.count_labels(h_rows_cl)
        torch.cuda.synchronize()
        tc = torch.tensor([time.perf_counter() - tw], device="cuda", dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(tc, op=dist.ReduceOp.MAX)
        cl = {"value": world * B * Kc / float(tc.item()), "unit": UNIT, "h2d_bytes_per_step": B * 8,
              "d2h_bytes_per_step": int(cnt.size) * 8, "steps": Kc,
              "api": "ggq_count_labels (host row ids in, per-column counters out; row sets never leave the device)",
              "labels_counted_per_step": int(cnt.sum())}

    # ---- C5 (BASELINE.json configs[4]): the column-sharded configuration, a collective over all ranks; needs the
    # memory of the C4 replica, so everything that touches `gm` is read out first
    device_image_gb = gm.device_bytes / 1e9
    ws.close()
    gm.close()
    c5 = None
    if dist is not None and world >= 2 and not args.no_c5:
        torch.cuda.empty_cache()
        c5 = c5_measure(args, dist, rank, world, local, 10, 2)

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the descent kernel (8 + 8 + 8 k_t + 32 V_t bytes per row) ---
    peak, peak_src = peaks()
    k_ms = float(np.mean(kern_ms))
    alg_bytes = (16.0 * B * K + 8.0 * ids + 32.0 * touches) / K
    # what this layout must move: 64-byte sibling blocks + one 32-byte sector per 1-byte leaf-code probe
    layout_bytes = (8.0 * B * K + 4.0 * ids + 64.0 * probes + 32.0 * code_probes) / K
    achieved = alg_bytes / (k_ms / 1e3) / 1e9
    roof = {"bound": "hbm", "kernel": "brwt_rows_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak, "traffic": None, "peak_source": peak_src, "kernel_ms": k_ms,
            "algorithmic_bytes_per_launch": alg_bytes,
            "bytes_per_row": alg_bytes / B, "node_touches_per_row": touches / (B * K),
            "layout_bytes_per_launch": layout_bytes, "layout_achieved": layout_bytes / (k_ms / 1e3) / 1e9,
            "layout_frac": layout_bytes / (k_ms / 1e3) / 1e9 / peak,
            "kernel_share_of_step": k_ms / ms_step,
            "probes_per_row": probes / (B * K), "leaf_code_probes_per_row": code_probes / (B * K),
            "leaf_code_escapes_per_row": code_esc / (B * K),
            "touches_note": "node_touches_per_row is V_t of the reference algorithm (SURVEY.md 8d), counted by the plain "
                            "descent on the same batches; the coded descent performs fewer"}
    if plain_ms:
        roof["plain_descent_kernel_ms"] = float(np.mean(plain_ms))
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        try:
            roof["traffic"] = json.load(open(tr)).get("brwt_rows_kernel_dram_bytes_per_launch")
        except Exception:
            pass

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic", "config": workload_config(args, world),
            "clocks": clk, "roofline": roof, "gpu_launches": launches,
            "build_seconds": build_s, "device_image_gb": device_image_gb, "leaf_codes": codes,
            "column_ids_per_row": ids / (B * K)}
    line["spot_check"] = spot
    if extra is not None:
        line["other_calls_host_to_host"] = extra
    if e2e is not None:
        line["e2e"] = e2e
        line["e2e_u32"] = e2e_u32
        line["e2e_u64"] = e2e_u64
    if cl is not None:
        line["count_labels_e2e"] = cl
    if world == 1 and not args.no_cpu_baseline:
        built = build_cpu_reference(args)
        if built is not None:
            R, rm, sub_n, threads = built
            rate, cnt, secs = cpu_sample_rate(rm, sub_n, threads, 15.0)
            line["cpu_baseline"] = {
                "value": rate, "unit": UNIT, "cores": threads, "kind": "reference",
                "sample": "%d random rows in %.1f s on a %d x %d d=%g sub-shape (same generator; the reference "
                          "builder cannot make the full shape); unmodified BRWT::get_row under omp parallel for" %
                          (cnt, secs, sub_n, args.cols, args.density)}
    if c5 is not None:
        line["c5"] = c5
    # ---- the metric's "rank/select ops/s" term (BASELINE.json configs[2]): independent vectors =>
    # replicas only, measured on rank 0
    if not args.no_bitvector:
        line["bitvector"] = bitvector_block(cpu=(world == 1 and not args.no_cpu_baseline))
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------
# C5: RefSeq-shaped matrix, column-sharded across all ranks: the image (~180 GB in the slot layout)
# does not fit one GPU.  Every rank answers the SAME batch over its own subtrees; partial CSRs are exchanged with
# NCCL (all-gather of offsets + grouped broadcasts of column ids) and merged in shard order on every rank.
# --------------------------------------------------------------------------------------------
def d2h_array(ptr, nbytes, dtype):
    """Device pointer (int) -> numpy array (outside every timed region)."""
    from cuda.bindings import runtime as rt
    out = np.empty(nbytes // np.dtype(dtype).itemsize, dtype=dtype)
    if nbytes:
        (err,) = rt.cudaMemcpy(out.ctypes.data, ptr, nbytes, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)
        if int(err) != 0:
            raise RuntimeError("cudaMemcpy failed: %s" % err)
    return out


def c5_measure(args, dist, rank, world, local, K, W):
    """C5 on an initialised process group: builds this rank's shard, times K collective steps, checks a sample of the
    answers against direct evaluation.  Returns the result block on rank 0 (None elsewhere)."""
    import torch
    import genome_graph_annotation_b200 as g
    ids = [g.Comm.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0)
    comm = g.Comm(ids[0], rank, world, local)
    n, m, d, B = int(args.c5_rows), args.c5_cols, args.c5_density, args.c5_batch
    slice_mode = not args.c5_full
    clocks = ClockSampler(local)
    clocks.start()   # (well before the timed region: nvidia-smi takes a while to come up)
    t0 = time.time()
    gm = g.GpuMatrix.synthesize_brwt_shard(n, m, d, rank, world, seed=SEED, arity=args.arity, device=local)
    torch.cuda.synchronize()
    build_s = time.time() - t0
    ws = gm.workspace()
    stream = torch.cuda.ExternalStream(ws.stream, device=torch.device("cuda", local))
    gen = torch.Generator(device="cuda")
    n_pool = min(K + W, 4)
    batches = []
    for b in range(n_pool):
        gen.manual_seed(SEED + b)        # the same batch on every rank
        batches.append(torch.randint(0, n, (B,), device="cuda", dtype=torch.int64, generator=gen))

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    def step(i):
        """-> (first query, queries, d_offsets, d_cols, ids) of what this rank ends up holding"""
        if slice_mode:
            return ws.get_rows_sharded_slice_device(comm, batches[i % n_pool].data_ptr(), B, 8)
        off, cols, total = ws.get_rows_sharded_device(comm, batches[i % n_pool].data_ptr(), B, 8)
        return 0, B, off, cols, total

    if not args.no_peer:
        # merge straight from the peers' partial CSRs over NVLink (cudaIpc); capacity: this batch size, ids with slack
        comm.enable_peer(ws, B, int(B * m * d * 1.3 / world) + (1 << 20))
    for i in range(W):
        step(i)
    barrier()
    clocks.mark_begin()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    kern_ms, xch_ms, xch_bytes, touches, ids_total, launches = [], [], 0, 0, 0, 0
    with torch.cuda.stream(stream):
        e0.record()
    for i in range(K):
        total = step(W + i)[4]
        kern_ms.append(ws.stat("kernel_ms")); xch_ms.append(ws.stat("exchange_ms"))
        xch_bytes += int(ws.stat("exchange_bytes")); touches += int(ws.stat("touches"))
        launches += int(ws.stat("launches")); ids_total += total
    with torch.cuda.stream(stream):
        e1.record()
    barrier()
    clk = clocks.stop()
    # spot check (untimed): the first rows this rank holds of batch 0 against direct evaluation of the implicit matrix
    q0, nq, d_off, d_cols, total = step(0)
    torch.cuda.synchronize()
    ns = min(nq, args.c5_spot_rows)
    h_off = d2h_array(d_off, (ns + 1) * 8, np.uint64)
    h_cols = d2h_array(d_cols, int(h_off[ns]) * 8, np.uint64)
    sample = batches[0][q0:q0 + ns].cpu().numpy().view(np.uint64)
    counts, cols = direct_rows(sample, m, d)
    same = np.array_equal(np.diff(h_off.astype(np.int64)), counts.astype(np.int64)) and np.array_equal(h_cols, cols)
    ok = torch.tensor([1.0 if same else 0.0], device="cuda", dtype=torch.float64)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    if float(ok.item()) != 1.0:
        raise RuntimeError("bench.py: C5 rows differ from direct evaluation")
    t = torch.tensor([e0.elapsed_time(e1), float(np.mean(kern_ms)), float(np.mean(xch_ms)), float(touches),
                      float(gm.device_bytes), float(ids_total)], device="cuda", dtype=torch.float64)
    tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
    ms_step = float(tmax[0]) / K
    peer = comm.peer_active
    block = None
    if rank == 0:
        peak, peak_src = peaks()
        k_ms = float(tmax[1])
        ids_all = float(tsum[5]) if slice_mode else float(ids_total)   # ids of the whole batch over the K steps
        # the slowest rank's descent over ITS subtrees bounds the step: algorithmic bytes of all ranks / that time
        alg = (16.0 * B * K * world + 8.0 * ids_all + 32.0 * float(tsum[3])) / K
        how = ("merge kernels read the shards' partial rows from peer memory over NVLink (cudaIpc)" if peer
               else "NCCL exchange + merge")
        block = {"value": B / (ms_step / 1e3), "unit": UNIT, "ms_per_step": ms_step, "steps": K, "warmup": W, "scaling": "strong",
                 "config": {"workload": "C5 RefSeq-shaped synthetic BRWT %.3g x %d density %g, arity-%d basic partitioner, "
                                        "column-sharded over %d GPUs; get_rows over %d uniformly random row ids per step "
                                        "(same batch on every rank), %s, %s" %
                                        (n, m, d, args.arity, world, B, how,
                                         "every rank merges its 1/%d slice of the batch" % world if slice_mode
                                         else "full CSR on every rank"),
                            "rows": n, "cols": m, "density": d, "arity": args.arity, "batch_rows": B, "seed": SEED,
                            "result": "slice per rank" if slice_mode else "full CSR on every rank",
                            "parallelism": "column shards x%d (cut frontier dealt in DFS runs), %s" %
                                           (world, "peer-memory merge (P2P loads), 8-byte all-gathers as barriers" if peer
                                            else "ncclAllGather + grouped ncclBroadcast"),
                            "l2": "inputs larger than L2"},
                 "clocks": clk,
                 "roofline": {"bound": "hbm", "kernel": "brwt_rows_kernel", "achieved": alg / (k_ms / 1e3) / 1e9 / world,
                              "peak": peak, "unit": "GB/s", "frac": alg / (k_ms / 1e3) / 1e9 / world / peak, "traffic": None,
                              "peak_source": peak_src, "kernel_ms_max_over_ranks": k_ms,
                              "note": "per-GPU average: algorithmic bytes of all shards / (ranks x slowest descent)"},
                 "exchange": {"mode": "peer (timed: fused transfer + merge)" if peer else "nccl (timed: transfer only)",
                              "ms_max_over_ranks": float(tmax[2]), "bytes_per_step_per_rank": xch_bytes / K,
                              "gb_per_s_per_rank": xch_bytes / K / (float(tmax[2]) / 1e3) / 1e9 if float(tmax[2]) > 0 else None},
                 "spot_check": {"rows_per_rank": int(ns), "identical": True,
                                "against": "direct numpy evaluation of the implicit matrix, counts and ids in DFS order"},
                 "gpu_launches": launches, "build_seconds": build_s,
                 "device_image_gb_per_rank_max": float(tmax[4]) / 1e9, "device_image_gb_total": float(tsum[4]) / 1e9,
                 "column_ids_per_row": ids_all / (B * K)}
    ws.close(); gm.close(); comm.close()
    return block


def run_c5(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world < 2:
        raise RuntimeError("--workload c5 is the column-sharded configuration: launch with torchrun, >= 2 ranks")
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    block = c5_measure(args, dist, rank, world, local, args.steps, args.warmup)
    if rank == 0:
        line = {"metric": METRIC, "n_gpus": world, "higher_is_better": True, "vs_baseline": None, "dtype": "u64",
                "data": "synthetic"}
        line.update(block)
        emit(line)
    dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    quiet_stdout()
    if a.impl == "reference":
        run_reference(a)
    elif a.workload == "c5":
        run_c5(a)
    else:
        run_ours(a)
