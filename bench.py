#!/usr/bin/env python
"""bench.py -- projected alignments/sec of the genome->transcriptome projection hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--names M]

A "step" is one pass of the whole kernel sequence (lookup -> chain walk -> pairing -> CIGAR merge)
over one synthetic bundle.  Workload = BASELINE.json configs[1] shape (C2): GENCODE-shaped
annotation (~250k transcripts, ~2.2M exons, 25 chromosomes) + stranded (--fr) paired-end 150 bp
name-collated reads, `--names` read names per GPU (default 2M names = ~4.6M alignments; the full C2
is 100M pairs, scaled down so that the default run finishes in minutes -- see DESIGN.md).
Ranks shard by read name (weak scaling: every GPU gets its own seeded shard, annotation replicated,
no collective on the data path).

  value  : alignments/s, inputs resident in HBM, CUDA-event time over K steps, max over ranks
  e2e    : alignments/s through brb_project() with pinned HOST columns (H2D + kernels + D2H per step)
  roofline / cpu_baseline : see DESIGN.md "Measurement"
"""
import argparse
import re
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def _peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region: NVML polled every millisecond (the timed region
    of the resident leg is ~10 ms, shorter than one nvidia-smi call), nvidia-smi as the fallback."""

    def __init__(self, gpu):
        self.gpu, self.rows, self.stop_flag, self.th, self.how = gpu, [], False, None, None

    def _nvml_open(self):
        import pynvml as nv
        nv.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ent = vis.split(",")[self.gpu].strip()
            h = nv.nvmlDeviceGetHandleByIndex(int(ent)) if ent.isdigit() else nv.nvmlDeviceGetHandleByUUID(ent)
        else:
            h = nv.nvmlDeviceGetHandleByIndex(self.gpu)
        self.nv, self.h, self.mx = nv, h, nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        self._nvml_sample()              # the first call pays the set-up; it is made before the timed region starts
        self.rows.clear()
        self.how = "nvml, polled back to back"

    def _nvml_sample(self):
        nv, h = self.nv, self.h
        bits = [(nv.nvmlClocksEventReasonHwSlowdown, 0), (nv.nvmlClocksEventReasonHwThermalSlowdown, 1),
                (nv.nvmlClocksEventReasonSwThermalSlowdown, 2), (nv.nvmlClocksEventReasonSwPowerCap, 3)]
        r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
        row = [str(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), str(self.mx), "", "", "", ""]
        for b, i in bits:
            if r & b:
                row[2 + i] = "active"
        self.rows.append(row)

    def _nvml(self):
        while not self.stop_flag:
            self._nvml_sample()
            time.sleep(0.0005)

    def _smi(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        self.how = "nvidia-smi, 0.1 s period"
        while not self.stop_flag:
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([x.strip() for x in o.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def prepare(self):
        """Open NVML (before the barrier that starts the timed region); nvidia-smi is used when that fails."""
        try:
            self._nvml_open()
        except Exception:
            self.nv = None

    def _run(self):
        if getattr(self, "nv", None) is not None:
            try:
                self._nvml()
                return
            except Exception:
                pass
        self._smi()

    def start(self):
        self.th = threading.Thread(target=self._run, daemon=True)
        self.th.start()

    def stop(self):
        self.stop_flag = True
        if self.th:
            self.th.join(timeout=6)
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        mx = max([int(r[1]) for r in self.rows if r[1].isdigit()] or [0])
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if len(r) >= 6 and r[2 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": reasons, "samples": len(sm),
                "how": self.how}


def _traffic(names):
    """dram__bytes_read.sum + dram__bytes_write.sum per step from the newest committed `ncu --set full` capture
    (profiles/r*_traffic.json), valid for the default workload size only; else None."""
    try:
        if names != 2_000_000:
            return None, None, None
        import glob
        import re
        f = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")), key=lambda x: [int(t) for t in re.findall(r"\d+", os.path.basename(x))])[-1]
        with open(f) as fh:
            t = json.load(fh)
        return int(sum(t.values())), {k: int(v) for k, v in t.items()}, os.path.relpath(f, ROOT)
    except Exception:
        return None, None, None


def workload(seed, names, kind="c2"):
    """c2 (default, the configuration the metric is quoted on): stranded paired-end reads.  c5: the multi-mapping stress
    shape of BASELINE.json configs[4] -- unstranded (both strand tables searched), 1 + Geometric(5) fragments per read
    name capped at 50 (many secondary alignments, large name groups)."""
    from bramble_b200 import synth_fast as sf
    A = sf.make_annotation(2002)
    ix = sf.build_index(A)
    if kind == "c5":
        B = sf.make_short_batch(A, ix, seed, names, stranded=None, multi_mean=5.0)
    else:
        B = sf.make_short_batch(A, ix, seed, names, stranded="fr")
    return A, ix, B


WORKLOAD_FLAGS = {"c2": ["--fr"], "c5": []}


def ref_cmd(cores, kind="c2"):
    return WORKLOAD_FLAGS[kind] + ["-p", str(max(1, min(cores, 255)))]


def time_reference(A, ix, B, n_small, n_big, cores, repeats=1, binary=None, extra=(), kind="c2"):
    """Reference C++ (oracle/_ref/bramble_ref, unmodified sources) on a bounded sample of the same
    workload.  Throughput is MARGINAL: (aln_big - aln_small) / (t_big - t_small), which removes the GTF
    load + index build the reference pays once per run.  BGZF level 1 for the output, all host cores."""
    from bramble_b200 import synth, synth_fast as sf
    ref = binary or os.path.join(ROOT, "oracle", "_ref", "bramble_ref")
    if not os.path.exists(ref):
        return None
    d = tempfile.mkdtemp(prefix="brb_bench_")
    sf.write_gtf(A, os.path.join(d, "a.gtf"))

    class Ann:
        chroms, chrom_len = A.chroms, A.chrom_len
    out = []
    pipe = {}                 # bramble_b200_cli only: its own clock from the first input block to the closed output (BRB_CLI_TIMING)
    for n in (n_small, n_big):
        recs = sf.to_records(B, 0, n)
        bam = os.path.join(d, f"in{n}.bam")
        synth.write_bam(bam, Ann, recs)
        n_aln = len(recs)
        best = None
        for _ in range(repeats):
            wd = tempfile.mkdtemp(dir=d)
            env = dict(os.environ, BRB_SHIM_LEVEL="1", BRB_CLI_TIMING="1")     # (the second one only means something to bramble_b200_cli)
            t0 = time.perf_counter()
            r = subprocess.run([ref, bam, "-G", os.path.join(d, "a.gtf"), "-o", os.path.join(wd, "out.bam"), "--quiet"] + ref_cmd(cores, kind) + list(extra),
                               cwd=wd, env=env, capture_output=True, text=True)
            dt = time.perf_counter() - t0
            if r.returncode != 0:
                raise RuntimeError("reference failed: " + r.stderr[-500:])
            best = dt if best is None else min(best, dt)
            m = re.search(r"\[timing\] pipeline ([0-9.]+)s", r.stderr or "")
            if m:
                pipe[n] = min(pipe.get(n, 1e30), float(m.group(1)))
        out.append((n_aln, best))
    (a0, t0), (a1, t1) = out
    if t1 - t0 < 0.05:
        return None                                   # the sample is too small to resolve a marginal rate
    rate = (a1 - a0) / (t1 - t0)
    return dict(value=rate, n_aln=a1 - a0, seconds=t1 - t0, startup_s=t0, n_aln_big=a1, pipeline_s=pipe.get(n_big), sample=f"first {n_big} read names ({a1} alignments) of the rank-0 shard, "
                f"marginal over a {n_small}-name run (startup {t0:.1f}s excluded), BAM in -> BAM out, BGZF level 1")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--names", type=int, default=2_000_000)
    ap.add_argument("--ref-names", type=int, default=150_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="c2", choices=["c2", "c5"], help="c2: the metric's configuration (default); c5: multi-mapping stress shape")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs: resident steps only (also skips the ordered-rows leg)")
    ap.add_argument("--e2e-bundles", type=int, default=8, help="bundles per step on the end-to-end path")
    ap.add_argument("--e2e-rows", default="packed", choices=["packed", "columns"],
                    help="row layout on the end-to-end path: BRB_OPT_PACKED_ROWS (16-byte records, fewer bytes device->host) or the plain columns")
    ap.add_argument("--no-bind", action="store_true", help="N>1: do not pin each rank to the CPUs of its GPU's NUMA node")
    ap.add_argument("--e2e-contexts", type=int, default=3, help="brb_ctx (host threads / streams) on the end-to-end path")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cores = os.cpu_count() or 1
    unit = "alignments/s"
    metric = "projected alignments/sec (input BAM records consumed)"
    if args.workload == "c5" and args.names == 2_000_000:
        args.names = 400_000                  # ~10 alignments per name: about the same alignments per GPU as c2
    if args.workload == "c5" and args.ref_names == 150_000:
        args.ref_names = 30_000
    shape = {"c2": "C2: synthetic GENCODE-shaped GTF (252,687 transcripts, 2.24M exons, 25 chromosomes) + --fr paired-end 150bp name-collated reads",
             "c5": "C5 shape: the same annotation + unstranded paired-end 150bp reads, 1+Geometric(5) fragments per read name (cap 50)"}[args.workload]
    config = {"workload": f"{shape}, {args.names} read names per GPU", "flags": " ".join(WORKLOAD_FLAGS[args.workload]) or "(default)", "names_per_gpu": args.names,
              "l2": "inputs + intermediates (>1 GB per step) exceed the 126 MB L2"}

    if args.impl == "reference":
        if rank != 0:
            return
        A, ix, B = workload(1000, max(args.ref_names, 2000), args.workload)
        vals = []
        res = None
        for _ in range(max(1, min(args.steps, 3))):
            res = time_reference(A, ix, B, 2000, args.ref_names, cores, kind=args.workload)
            if res is None:
                print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/bramble_ref not built (reference sources absent)"}))
                return
            vals.append(res["value"])
        v = max(vals)
        print(json.dumps({"impl": "reference", "metric": metric, "value": v, "unit": unit, "n_gpus": args.gpus, "steps": args.steps,
                          "warmup": args.warmup, "ms_per_step": 1e3 * res["seconds"], "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": v, "unit": unit, "cores": min(cores, 255), "kind": "reference", "sample": res["sample"]},
                          "e2e": {"value": v, "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    # stdout carries exactly one JSON line: whatever libraries print there (NCCL's version banner under torchrun) goes to
    # stderr instead; the line itself is written to the saved descriptor at the end
    sys.stdout.flush()
    out_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.write(out_fd, (json.dumps(line) + "\n").encode())

    import torch
    import torch.distributed as dist
    from bramble_b200 import capi, engine
    import __graft_entry__ as g
    g.build_engine()
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    from bramble_b200 import hostbind
    # one process per GPU: keep the rank's threads and its pinned staging pages on the GPU's NUMA node
    host = hostbind.bind_to_gpu_node(local, local, int(os.environ.get("LOCAL_WORLD_SIZE", world))) if (world > 1 and not args.no_bind) else hostbind.describe(local)
    host["cpus"] = os.cpu_count()
    A, ix, B = workload(1000 + rank, args.names, args.workload)
    n_aln = int(B.desc.n_aln)
    p = capi.resolve_params()            # --fr only affects ingest (strand column), thresholds are the short-read preset
    eng = engine.Engine(ix, device=local)
    tbar = torch.zeros(1, device=f"cuda:{local}")

    def barrier():
        if world > 1:
            dist.all_reduce(tbar)
        torch.cuda.synchronize()

    # ---- resident (kernel-only) timing
    eng.upload(B)
    # algorithmic bytes are defined on the reference's algorithm, which looks up EVERY read exon: take
    # them once, untimed, from the generic front kernel; the timed path is the engine's default choice
    checks = {}
    eng.set_option(capi.OPT_GENERIC_FRONT, 1)
    eng.run_resident(p)
    alg_ref, alg_front = int(eng.stats().alg_bytes), int(eng.stats().alg_bytes_front)
    if not args.no_e2e:       # full-size parity property: every engine variant must produce the same row set (order-independent checksum)
        checks["generic_front"] = engine.rows_checksum(eng.download(B, raw=True))
    eng.set_option(capi.OPT_GENERIC_FRONT, 0)
    for _ in range(max(3, args.warmup)):
        eng.run_resident(p)
    st = eng.stats()
    sampler = ClockSampler(local)
    sampler.prepare()                     # every rank watches its own GPU; rank 0's samples go into `clocks`, the medians of all into `host`
    import gc
    gc.collect()
    gc.disable()                          # a collection pass between two steps is a millisecond of idle GPU in an 11 ms timed region
    barrier()
    sampler.start()
    phase = {k: 0.0 for k in ("ms_lookup", "ms_chain", "ms_rescue", "ms_finalize", "ms_pair", "ms_merge")}
    eng.timer_start()
    launches = 0
    per_step = []
    for _ in range(args.steps):           # nothing but the call and one struct copy between steps: the GPU idles while the host is busy
        eng.run_resident(p)
        per_step.append(eng.stats())
    ms = eng.timer_stop()
    gc.enable()
    for s in per_step:
        launches += int(s.n_launches)
        for k in phase:
            phase[k] += getattr(s, k)
    barrier()
    clocks = sampler.stop()
    # the same with BRB_OPT_ORDERED_ROWS: one more streaming pass (scan + k_place) lays the rows out in emit order on the device
    ms_ordered = ms_packed = None
    if not args.no_e2e:
        eng.set_option(capi.OPT_ORDERED_ROWS, 1)
        for _ in range(3):
            eng.run_resident(p)
        eng.timer_start()
        for _ in range(args.steps):
            eng.run_resident(p)
        ms_ordered = eng.timer_stop()
        checks["ordered_rows"] = engine.rows_checksum(eng.download(B, raw=True))
        eng.set_option(capi.OPT_ORDERED_ROWS, 0)
        eng.run_resident(p)
        checks["default"] = engine.rows_checksum(eng.download(B, raw=True))
        eng.set_option(capi.OPT_PACKED_ROWS, 1)
        for _ in range(3):
            eng.run_resident(p)
        eng.timer_start()
        for _ in range(args.steps):
            eng.run_resident(p)
        ms_packed = eng.timer_stop()
        checks["packed_rows"] = engine.rows_checksum(eng.download(B, raw=True))
        eng.set_option(capi.OPT_PACKED_ROWS, 0)
    barrier()
    st = eng.stats()
    # ---- end to end through the C ABI with pinned host columns: the step's alignments as `--e2e-bundles` bundles
    # (cut at read-name boundaries, like the reference's bundles) through `--e2e-contexts` contexts, one host
    # thread each, so that the H2D copy of a bundle overlaps the kernels of another and the D2H copy of a third;
    # every brb_project() call = H2D of its bundle + all kernels + D2H of its rows
    PB, PBs, ctxs, nbun = None, [], [], 0
    e2e_s, e2e_single_s, e2e_h2d, e2e_d2h, e2e_rec = 0.0, 1.0, 0, 0, 0
    if not args.no_e2e:
        G = int(B.desc.n_groups)
        nbun = max(1, min(args.e2e_bundles, G))
        cuts = [G * i // nbun for i in range(nbun + 1)]
        PBs = [eng.pinned_copy(B.slice_groups(g0, g1)) for g0, g1 in zip(cuts[:-1], cuts[1:])]
        ctxs = [eng.context() for _ in range(max(1, args.e2e_contexts))]
        e2e_packed = args.e2e_rows == "packed"
        for c in ctxs:
            c.set_option(capi.OPT_PACKED_ROWS, int(e2e_packed))
        eng.set_option(capi.OPT_PACKED_ROWS, int(e2e_packed))
        engine.project_pipelined(ctxs, p, PBs)
        res = engine.project_pipelined(ctxs, p, PBs, checksum=True)
        checks["e2e_bundles"] = (sum(r[4] for r in res) & 0xFFFFFFFFFFFFFFFF, sum(r[0] for r in res))
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            res = engine.project_pipelined(ctxs, p, PBs)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        e2e_h2d, e2e_d2h, e2e_rec = sum(r[2] for r in res), sum(r[3] for r in res), sum(r[0] for r in res)
        barrier()
        # the same through ONE call on one context (no overlap), for comparison
        PB = eng.pinned_copy(B)
        eng.project(p, PB, raw=True)
        t0 = time.perf_counter()
        for _ in range(min(3, args.steps)):
            eng.project(p, PB, raw=True)
        torch.cuda.synchronize()
        e2e_single_s = (time.perf_counter() - t0) / min(3, args.steps)
        eng.set_option(capi.OPT_PACKED_ROWS, 0)
        barrier()
    t = torch.tensor([ms, e2e_s * 1e3, float(n_aln)], dtype=torch.float64, device=f"cuda:{local}")
    tmax = t.clone()
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    total_aln = float(t[2].item()) if world > 1 else float(n_aln)
    by_rank = [torch.zeros(6, dtype=torch.float64, device=f"cuda:{local}") for _ in range(world)]
    mine = torch.tensor([e2e_s * 1e3 / max(1, args.steps), float(host["numa_node"] if host["numa_node"] is not None else -1),
                         float(host["cpus_allowed"]), ms / max(1, args.steps), float(clocks["sm_mhz"] or 0),
                         float(any(r != "sw_power_cap" for r in clocks["reasons"]))], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_gather(by_rank, mine)
    else:
        by_rank = [mine]
    ms_max, e2e_ms_max = float(tmax[0].item()), float(tmax[1].item())
    if rank == 0:
        peak, peak_src = _peaks()
        step_ms = ms_max / args.steps
        value = total_aln * args.steps / (ms_max / 1e3)
        e2e = total_aln * args.steps / (e2e_ms_max / 1e3) if not args.no_e2e else None
        alg = alg_ref
        achieved = alg / (ms / args.steps / 1e3) / 1e9
        names = {"ms_lookup": "k_front_fast (lookup + evaluate)", "ms_chain": "k_chain (rescue path only)", "ms_finalize": "k_build (rescue path only)",
                 "ms_pair": "k_back (pairing + CIGAR merge + rows)", "ms_merge": "scan + k_place + k_group_rows (BRB_OPT_ORDERED_ROWS only)",
                 "ms_rescue": "k_ksw (rescue path only)"}
        top = max(phase, key=lambda k: phase[k])
        traffic, traffic_by_kernel, traffic_src = _traffic(args.names if args.workload == "c2" else -1)
        alg_of = {"ms_lookup": alg_front, "ms_pair": alg - alg_front, "ms_merge": alg - alg_front}
        top_ms = phase[top] / args.steps
        top_alg = alg_of.get(top, alg)
        line = {"metric": metric, "value": value, "unit": unit, "n_gpus": args.gpus, "steps": args.steps, "warmup": max(3, args.warmup),
                "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
                "data": "synthetic", "config": config, "gpu_launches": launches, "clocks": clocks,
                "e2e": None if args.no_e2e else {"value": e2e, "unit": unit, "h2d_bytes_per_step": e2e_h2d, "d2h_bytes_per_step": e2e_d2h,
                        "records_per_step": e2e_rec, "bundles_per_step": nbun, "contexts": len(ctxs),
                        "how": "brb_project() per bundle from pinned host columns, one host thread per context",
                        "rows": "BRB_OPT_PACKED_ROWS: one 16-byte record per row" if args.e2e_rows == "packed" else "columns (22 bytes per row + CIGAR CSR)",
                        "single_call_value": n_aln / e2e_single_s, "single_call_note": "rank 0, one brb_project() call for the whole step, no overlap"},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                             "traffic_by_kernel": traffic_by_kernel, "traffic_source": f"{traffic_src} (ncu --set full, one step)" if traffic_src else None,
                             "peak_source": peak_src, "scope": "whole kernel sequence of one step (algorithmic bytes of SURVEY 8(d) / DESIGN.md)",
                             "alg_bytes_per_step": alg, "alg_bytes_per_alignment": alg / max(1, n_aln),
                             "top_kernel": names[top], "top_kernel_share": phase[top] / max(1e-9, sum(phase.values())),
                             "top_kernel_ms": top_ms, "top_kernel_alg_bytes": top_alg,
                             "top_kernel_achieved": top_alg / (top_ms / 1e3) / 1e9, "top_kernel_frac": top_alg / (top_ms / 1e3) / 1e9 / peak,
                             "phase_ms_per_step": {names[k]: v / args.steps for k, v in phase.items()},
                             "ordered_rows": None if ms_ordered is None else {"ms_per_step": ms_ordered / args.steps, "frac": alg / (ms_ordered / args.steps / 1e3) / 1e9 / peak,
                                              "what": "rank 0, BRB_OPT_ORDERED_ROWS=1: rows moved into emit order on the device (scan + k_place + k_group_rows) "
                                                      "instead of unit blocks + aln_rec_off/aln_rec_cnt"},
                             "packed_rows": None if ms_packed is None else {"ms_per_step": ms_packed / args.steps, "frac": alg / (ms_packed / args.steps / 1e3) / 1e9 / peak,
                                              "what": "rank 0, BRB_OPT_PACKED_ROWS=1: the same rows written as 16-byte records (same algorithmic bytes: they are "
                                                      "defined on the reference's row content, not on this layout)"}},
                "parity_at_bench_size": None if not checks else {
                    "property": "order-independent checksum of all rows (source alignment, tid, pos, strand, flags, CIGAR) and row count agree between "
                                "the default kernels, the generic front kernel, BRB_OPT_ORDERED_ROWS, BRB_OPT_PACKED_ROWS and the 8-bundle / 3-context end-to-end path (rank 0)",
                    "rows": checks["default"][1], "checksum": "%016x" % checks["default"][0],
                    "agree": len({(int(c[0]), int(c[1])) for c in checks.values()}) == 1},
                "host": {**host, "e2e_ms_per_step_by_rank": [round(float(x[0].item()), 3) for x in by_rank],
                         "numa_node_by_rank": [int(x[1].item()) for x in by_rank], "cpus_allowed_by_rank": [int(x[2].item()) for x in by_rank],
                         "ms_per_step_by_rank": [round(float(x[3].item()), 4) for x in by_rank], "sm_mhz_by_rank": [int(x[4].item()) for x in by_rank],
                         "slowdown_seen_by_rank": [bool(x[5].item()) for x in by_rank]},
                "work": {"alignments_per_gpu": n_aln, "candidates": int(st.n_candidates), "matches": int(st.n_matches), "records": int(st.n_rec)}}
        if args.gpus == 1 and not args.no_cpu_baseline:
            try:
                res = time_reference(A, ix, B, 2000, args.ref_names, cores, kind=args.workload)
            except Exception as e:     # the baseline is a reported number, never a reason to lose the bench line
                res = None
                line["cpu_baseline_error"] = str(e)[:200]
            if res:
                line["cpu_baseline"] = {"value": res["value"], "unit": unit, "cores": min(cores, 255), "kind": "reference", "sample": res["sample"]}
            # the drop-in CLI (host C++: BGZF/BAM decode, ingest, flush replay, record assembly, BGZF encode; CUDA engine
            # through the C ABI) on the same BAM sample, same flags: end to end INCLUDING BGZF, marginal rate like the reference
            try:
                cli_names = min(3 * args.ref_names, int(B.desc.n_groups))      # a larger sample: the CLI needs several bundles in flight
                cli = time_reference(A, ix, B, 2000, cli_names, cores, repeats=3, binary=g.build_cli(), extra=["--level", "1"], kind=args.workload)   # best of 3 each: start-up (CUDA context) varies by tenths of a second
                if cli and cli.get("pipeline_s"):
                    # the CLI's own clock (input open -> output closed, index build and CUDA start-up excluded) resolves a
                    # sub-second pipeline better than the difference of two process wall times
                    line["cli_e2e"] = {"value": cli["n_aln_big"] / cli["pipeline_s"], "unit": unit, "sample": cli["sample"], "host_threads": min(cores, 255),
                                       "pipeline_s": cli["pipeline_s"], "marginal_value": cli["value"],
                                       "what": "bramble_b200_cli BAM in -> BAM out (1 GPU): alignments of the sample / the CLI's pipeline time "
                                               "(BRB_CLI_TIMING: first input block to closed output), best of 3 runs; marginal_value = the same by "
                                               "the reference arm's method (difference of two process wall times)"}
                elif cli:
                    line["cli_e2e"] = {"value": cli["value"], "unit": unit, "sample": cli["sample"], "host_threads": min(cores, 255),
                                       "what": "bramble_b200_cli BAM in -> BAM out (1 GPU), marginal over a 2000-name run, best of 3 runs each"}
            except Exception as e:
                line["cli_e2e_error"] = str(e)[:200]
        emit(line)
    if PB is not None:
        eng.pinned_free(PB)
    for x in PBs:
        eng.pinned_free(x)
    for c in ctxs:
        c.close()
    eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
