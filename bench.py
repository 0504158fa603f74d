#!/usr/bin/env python
"""bench.py -- projected alignments/sec of the genome->transcriptome projection hot path.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--names M] [--workload c1..c5]

A "step" is one pass of the whole kernel sequence (lookup -> chain walk -> [clip rescue] -> pairing -> CIGAR merge)
over one synthetic bundle.  The headline workload is BASELINE.json configs[1] (C2): GENCODE-shaped annotation
(~250k transcripts, ~2.2M exons, 25 chromosomes) + stranded (--fr) paired-end 150 bp name-collated reads, `--names`
read names per GPU (default 2M names = ~4.6M alignments; the full C2 is 100M pairs, scaled down so that the default run
finishes in minutes -- see DESIGN.md).  Ranks shard by read name (weak scaling: every GPU gets its own seeded shard,
annotation replicated, no collective on the data path).

  value  : alignments/s, inputs resident in HBM, CUDA-event time over K steps, max over ranks
  e2e    : alignments/s through brb_project() with pinned HOST columns (H2D + kernels + D2H per step)
  roofline / cpu_baseline : see DESIGN.md "Measurement"
  workloads : (1 GPU only) the other four configurations -- c1, c3 (--lr -S, ksw2 clip rescue), c4 (--lr-hq --strict),
              c5 (paralog multi-mappers) -- each with kernel-only rate, per-kernel phase times, roofline on its own
              algorithmic bytes, an end-to-end call, the reference CPU with the matching flags, and bench-size parity
              against the CPU oracle (bramble_b200/workloads.py builds them).
"""
import argparse
import hashlib
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

UNIT = "alignments/s"
METRIC = "projected alignments/sec (input BAM records consumed)"
PHASES = ("ms_lookup", "ms_chain", "ms_rescue", "ms_finalize", "ms_pair", "ms_merge")


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


def engine_src_sha():
    """Identity of the kernels a profile was taken on: sha256 over the engine's device sources."""
    d = os.path.join(ROOT, "bramble_b200", "csrc")
    m = hashlib.sha256()
    for f in sorted(os.listdir(d)):
        if f.endswith((".cu", ".cuh", ".inl")):
            with open(os.path.join(d, f), "rb") as fh:
                m.update(f.encode() + b"\0" + fh.read())
    return m.hexdigest()[:16]


def _traffic(workload, names):
    """dram__bytes_read.sum + dram__bytes_write.sum per step from the newest committed `ncu --set full` capture
    (profiles/r*_traffic.json) -- only when that capture was taken on THESE kernel sources, this workload and this
    size (the file records all three); anything else is stale and reported as null."""
    try:
        import glob
        fs = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_traffic.json")), key=lambda x: [int(t) for t in re.findall(r"\d+", os.path.basename(x))])
        sha = engine_src_sha()
        for f in reversed(fs):
            with open(f) as fh:
                t = json.load(fh)
            if not isinstance(t.get("kernels"), dict):
                continue                                        # round-1 format: no provenance recorded
            if t.get("engine_src_sha") == sha and t.get("workload") == workload and int(t.get("names", -1)) == int(names):
                k = {a: int(b) for a, b in t["kernels"].items()}
                return int(sum(k.values())), k, os.path.relpath(f, ROOT)
    except Exception:
        pass
    return None, None, None


# ------------------------------------------------------------------------------------------------ reference CPU arm
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "bramble_ref")
SHIM_NOTE = ("the reference is the unmodified C++ compiled against oracle/shim (mini BGZF/BAM over zlib in place of htslib + libdeflate, "
             "which this image lacks): its BGZF share would be faster on a real htslib install, its projection would not change")


class RefBench:
    """The reference binary (or the drop-in CLI) on bounded samples of one workload.  Files are written once; a leg is a
    run on one sample and its rate is MARGINAL -- (aln_sample - aln_small) / (t_sample - t_small) -- which removes the
    GTF load + index build every run pays once (the small run is timed once per binary)."""

    def __init__(self, W, samples):
        """samples: {tag: read names}; must contain "small"."""
        from bramble_b200 import workloads
        self.W, self.d = W, tempfile.mkdtemp(prefix=f"brb_bench_{W.name}_")
        self.extra = workloads.write_annotation(W, self.d)
        self.n, self.t_small, self.decoy = {}, {}, False
        for tag, n in samples.items():
            self.n[tag] = (n, workloads.write_sample(W, os.path.join(self.d, f"{tag}.bam"), n))

    def run(self, tag, flags, env=None, binary=None, gtf="a.gtf", repeats=1):
        best, pipe = None, None
        for _ in range(repeats):
            wd = tempfile.mkdtemp(dir=self.d)
            e = dict(os.environ, BRB_SHIM_LEVEL="1", BRB_CLI_TIMING="1")
            e.update(env or {})
            t0 = time.perf_counter()
            r = subprocess.run([binary or REF_BIN, os.path.join(self.d, f"{tag}.bam"), "-G", os.path.join(self.d, gtf), "-o", os.path.join(wd, "out.bam"),
                                "--quiet"] + self.W.flags + self.extra + list(flags), cwd=wd, env=e, capture_output=True, text=True)
            dt = time.perf_counter() - t0
            if r.returncode != 0:
                raise RuntimeError("reference failed: " + (r.stderr or r.stdout)[-500:])
            best = dt if best is None else min(best, dt)
            m = re.search(r"\[timing\] pipeline ([0-9.]+)s", r.stderr or "")
            if m:
                pipe = float(m.group(1)) if pipe is None else min(pipe, float(m.group(1)))
            try:
                os.remove(os.path.join(wd, "out.bam"))
            except OSError:
                pass
        return best, pipe

    def leg(self, tag, flags, env=None, binary=None, gtf="a.gtf", repeats=1):
        """-> dict(value, seconds, n_aln, startup_s, pipeline_s) or None when the sample cannot resolve a marginal rate."""
        if gtf == "decoy.gtf" and not self.decoy:
            from bramble_b200 import workloads
            workloads.write_decoy_annotation(self.W, self.d)
            self.decoy = True
        key = (binary or REF_BIN, tuple(sorted((env or {}).items())) if binary else ())
        if key not in self.t_small:                 # start-up (GTF load, index build, CUDA context for the CLI): the same for every leg of a binary
            self.t_small[key] = self.run("small", flags, env, binary, "a.gtf", repeats)[0]
        t0 = self.t_small[key]
        t1, pipe = self.run(tag, flags, env, binary, gtf, repeats)
        a0, a1 = self.n["small"][1], self.n[tag][1]
        if t1 - t0 < 0.02:
            return None
        return dict(value=(a1 - a0) / (t1 - t0), seconds=t1 - t0, n_aln=a1 - a0, startup_s=t0, n_aln_big=a1, pipeline_s=pipe, t_big=t1)

    def sample_text(self, tag="big", what="BAM in -> BAM out, BGZF level 1"):
        return (f"first {self.n[tag][0]} read names ({self.n[tag][1]} alignments) of the rank-0 bundle, marginal over a "
                f"{self.n['small'][0]}-name run (start-up excluded), {what}")

    def cleanup(self):
        import shutil
        shutil.rmtree(self.d, ignore_errors=True)


def cpu_baseline(rb, cores, full_legs=False):
    """cpu_baseline object of one workload (rb: RefBench with samples "small", "big" and, for full_legs, "mid").  Always: the
    reference with all host cores, BAM -> BAM.  full_legs (headline workload): also `-p 1` (the deterministic run) and the
    projection-only time = `-p 1` with the shim's null writer minus the same run against a decoy annotation (decode +
    ingest only), SURVEY 8(d) / BASELINE.md section 3."""
    W = rb.W
    pN = max(1, min(cores, 255))
    main = rb.leg("big", ["-p", str(pN)])
    if main is None:
        return None
    out = {"value": main["value"], "unit": UNIT, "cores": pN, "kind": "reference", "sample": rb.sample_text(),
           "flags": " ".join(W.flags + (["-S genome.fa"] if W.genome is not None else [])) or "(default)", "seconds": main["seconds"], "note": SHIM_NOTE}
    if full_legs:
        l1 = rb.leg("mid", ["-p", "1"])
        ln = rb.leg("mid", ["-p", "1"], env={"BRB_SHIM_NULL_WRITE": "1"})
        ld = rb.leg("mid", ["-p", "1"], env={"BRB_SHIM_NULL_WRITE": "1"}, gtf="decoy.gtf")
        if l1:
            out["p1"] = {"value": l1["value"], "cores": 1, "what": "-p 1 (deterministic bundle order), BAM in -> BAM out", "sample": rb.sample_text("mid")}
        if ln and ld and ln["seconds"] > ld["seconds"]:
            out["projection_only"] = {
                "value": ln["n_aln"] / (ln["seconds"] - ld["seconds"]), "cores": 1, "unit": UNIT,
                "null_write_value": ln["value"], "decode_only_value": ld["value"],
                "what": "reference -p 1, output writer disabled (BRB_SHIM_NULL_WRITE=1), minus the same run against a decoy annotation "
                        "(renamed chromosomes: BAM decode + ingest + bundling, no projection) -- the CPU figure at the boundary of "
                        "`value` (convert_reads only)", "sample": rb.sample_text("mid", "no output")}
    return out


# ------------------------------------------------------------------------------------------------ GPU legs
def timed_resident(eng, p, steps):
    """K timed steps on the ctx stream (CUDA events), nothing but the call and one struct copy between steps."""
    per_step = []
    eng.timer_start()
    for _ in range(steps):
        eng.run_resident(p)
        per_step.append(eng.stats())
    ms = eng.timer_stop()
    phase = {k: sum(getattr(s, k) for s in per_step) / steps for k in PHASES}
    return ms, sum(int(s.n_launches) for s in per_step), phase


KERNEL_OF = {"short": {"ms_lookup": "k_front_fast (lookup + evaluate)", "ms_pair": "k_back (pairing + CIGAR merge + rows)",
                       "ms_merge": "scan + k_place + k_group_rows (BRB_OPT_ORDERED_ROWS only)"},
             "long": {"ms_lookup": "k_front (lookup + candidates [+ evaluate without -S])", "ms_chain": "k_chain (pass 1 of the chain walk; -S only)",
                      "ms_rescue": "clip rescue: plan + gather + k_ksw + finish (-S only)", "ms_finalize": "k_build (ideal CIGAR + scores; -S only)",
                      "ms_pair": "k_back (pairing + CIGAR merge + rows)", "ms_merge": "scan + k_place + k_group_rows (BRB_OPT_ORDERED_ROWS only)"}}


def oracle_parity(eng, W, min_names=200_000):
    """Bench-size parity against the CPU oracle: the first >= min_names read names of the bundle (all of it when it is
    smaller) through the engine and through oracle/libbrb_port.so; the order-independent row checksum (source alignment,
    tid, pos, strand, flags, similarity bits, clip score, CIGAR) and the row count must agree.  Untimed, except that the
    oracle reports the seconds of its own projection loop (one thread) -- the CPU figure at the engine's boundary."""
    import ctypes as C
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    from bramble_b200 import engine
    G = int(W.B.desc.n_groups)
    S = W.B if G <= min_names else W.B.slice_groups(0, min_names)
    got = engine.rows_checksum(eng.project(W.params, S, raw=True))
    o = oracle_lib.PortOut()
    if oracle_lib.port().brb_port_project(C.byref(W.ix.desc), C.byref(W.params), C.byref(S.desc), C.byref(o)) != 0:
        raise RuntimeError("oracle failed")
    want = engine.rows_checksum(o)
    secs, n_rescue = float(o.seconds_project), int(o.n_rescue)
    oracle_lib.port().brb_port_free(C.byref(o))
    n_aln = int(S.desc.n_aln)
    return {"oracle_agree": got == want, "names": min(G, min_names), "alignments": n_aln, "rows": got[1], "oracle_rows": want[1],
            "checksum": "%016x" % got[0], "oracle_checksum": "%016x" % want[0], "oracle_rescues": n_rescue,
            "oracle_port": {"value": n_aln / secs if secs > 0 else None, "unit": UNIT, "cores": 1, "kind": "port", "seconds": secs,
                            "what": "oracle/brb_port.cpp projection loop, one thread, columns in -> rows out (index build excluded)"}}


def measure_workload(W, args, local, cores, peak, with_cpu=True):
    """One of the non-headline configurations on one GPU: kernel-only rate + phases + roofline, an end-to-end call, parity
    against the oracle at bench size, the reference CPU with the matching flags."""
    from bramble_b200 import capi, engine
    eng = engine.Engine(W.ix, device=local)
    try:
        B, p = W.B, W.params
        n_aln = int(B.desc.n_aln)
        eng.upload(B)
        # reference-defined algorithmic bytes: every read exon looked up and every hit counted (BRB_OPT_GENERIC_FRONT), once, untimed
        eng.set_option(capi.OPT_GENERIC_FRONT, 1)
        eng.run_resident(p)
        alg, alg_front = int(eng.stats().alg_bytes), int(eng.stats().alg_bytes_front)
        eng.set_option(capi.OPT_GENERIC_FRONT, 0)
        for _ in range(3):
            eng.run_resident(p)
        st = eng.stats()
        steps = max(2, min(args.steps, int(1500.0 / max(st.kernel_ms, 0.05))))       # bounded: ~1.5 s of timed steps at most
        ms, launches, phase = timed_resident(eng, p, steps)
        st = eng.stats()
        step_ms = ms / steps
        names = KERNEL_OF["long" if W.long else "short"]
        top = max(phase, key=lambda k: phase[k])
        rec = {"config": {"workload": f"{W.shape}, {W.names} read names in one bundle", "flags": " ".join(W.flags + (["-S genome.fa"] if W.genome is not None else [])) or "(default)"},
               "value": n_aln / (step_ms / 1e3), "unit": UNIT, "ms_per_step": step_ms, "steps": steps, "gpu_launches_per_step": launches / steps,
               "work": {"alignments": n_aln, "candidates": int(st.n_candidates), "matches": int(st.n_matches), "records": int(st.n_rec),
                        "rescues": int(st.n_rescue), "cigar_words_in": int(B.arrays["cigar"].shape[0])},
               "roofline": {"bound": "hbm", "achieved": alg / (step_ms / 1e3) / 1e9, "peak": peak, "unit": "GB/s", "frac": alg / (step_ms / 1e3) / 1e9 / peak,
                            "basis": "algorithmic bytes (SURVEY 8(d)) of the whole kernel sequence / CUDA-event time of the sequence",
                            "traffic": None, "alg_bytes_per_step": alg, "alg_bytes_per_alignment": alg / max(1, n_aln),
                            "phase_ms_per_step": {names[k]: v for k, v in phase.items() if k in names and v > 0},
                            "top_kernel": names.get(top, top), "top_kernel_share": phase[top] / max(1e-9, sum(phase.values()))}}
        if W.name == "c3" and st.n_rescue:
            rec["roofline"]["rescue"] = {"problems_per_step": int(st.n_rescue), "ms_per_step": phase["ms_rescue"],
                                         "dp_cells_per_step": int(getattr(st, "ksw_cells", 0)) or None,
                                         "dp_cells_run_per_step": int(getattr(st, "ksw_cells_run", 0)) or None,
                                         "note": "dp_cells = query x target over all clip-rescue problems (what the reference computes); dp_cells_run = after "
                                                 "sharing the DP between identical problems of one read (same clip, same neighbour exons)"}
            if rec["roofline"]["rescue"]["dp_cells_per_step"]:
                rec["roofline"]["rescue"]["cell_updates_per_s"] = rec["roofline"]["rescue"]["dp_cells_per_step"] / max(1e-9, phase["ms_rescue"] / 1e3)
        if args.no_e2e:
            return rec
        # end to end: one brb_project() call from pinned host columns (H2D + kernels + D2H)
        want_rows = engine.rows_checksum(eng.download(B, raw=True))          # (the resident result of the timed steps)
        # compact input columns (16-bit counts / links, position instead of exon arrays) for short reads; a long-read bundle is CIGAR words
        # (90 % of its bytes), the compact forms save nothing there and cost an extra pass (measured: c4 40 -> 33 M alignments/s)
        PB = eng.pinned_copy(B if W.long else B.compact())
        try:
            eng.set_option(capi.OPT_PACKED_ROWS, 0 if W.long else 3)
            got_rows = engine.rows_checksum(eng.project(p, PB, raw=True), batch=PB)
            t0 = time.perf_counter()
            reps = 3
            for _ in range(reps):
                eng.project(p, PB, raw=True)
            dt = (time.perf_counter() - t0) / reps
            s2 = eng.stats()
            rec["e2e"] = {"value": n_aln / dt, "unit": UNIT, "h2d_bytes_per_step": int(s2.h2d_bytes), "d2h_bytes_per_step": int(s2.d2h_bytes),
                          "how": "one brb_project() call for the bundle from pinned host columns (short reads: compact input columns, rows as 8-byte records), no overlap between calls",
                          "rows_agree_with_resident_run": got_rows == want_rows}
            eng.set_option(capi.OPT_PACKED_ROWS, 0)
        finally:
            eng.pinned_free(PB)
        rec["parity_at_bench_size"] = oracle_parity(eng, W)
    finally:
        eng.close()
    if with_cpu:
        try:
            small, big = {"c1": (2000, 60_000), "c3": (500, 16_000), "c4": (500, 9_000), "c5": (1000, 24_000)}[W.name]
            rb = RefBench(W, {"small": small, "big": big})
            try:
                cb = cpu_baseline(rb, cores)
            finally:
                rb.cleanup()
            if cb:
                rec["cpu_baseline"] = cb
        except Exception as e:
            rec["cpu_baseline_error"] = str(e)[:300]
    return rec


def cli_e2e(rb, cores, gpus_list, repeats=2):
    """The drop-in CLI (host C++: BGZF/BAM decode, ingest, flush replay, record assembly, BGZF encode; CUDA engine through the
    C ABI), BAM in -> BAM out INCLUDING BGZF, the "cli" sample of rb and the flags of the reference arm, for every GPU count
    in gpus_list (the CLI's own dealer: bundle k -> device k mod N, ordered merge)."""
    import __graft_entry__ as g
    binary = g.build_cli()
    out = {}
    for n in gpus_list:
        leg = rb.leg("cli", ["-p", str(max(1, min(cores, 255))), "--level", "1"], env={"BRB_GPUS": str(n)}, binary=binary, repeats=repeats)
        if not leg:                         # the difference of two wall times did not resolve (start-up of N engines is most of both): the pipeline clock still stands
            _, pipe = rb.run("cli", ["-p", str(max(1, min(cores, 255))), "--level", "1"], {"BRB_GPUS": str(n)}, binary, "a.gtf", repeats)
            if not pipe:
                continue
            leg = {"value": None, "pipeline_s": pipe, "n_aln_big": rb.n["cli"][1]}
        r = {"unit": UNIT, "gpus": n, "host_threads": max(1, min(cores, 255)), "sample": rb.sample_text("cli"), "marginal_value": leg["value"]}
        if leg.get("pipeline_s"):
            r.update(value=leg["n_aln_big"] / leg["pipeline_s"], pipeline_s=leg["pipeline_s"])
        else:
            r.update(value=leg["value"])
        out[str(n)] = r
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--names", type=int, default=None, help="read names per GPU of the headline workload (default: the workload's bench size)")
    ap.add_argument("--ref-names", type=int, default=150_000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="c2", choices=["c1", "c2", "c3", "c4", "c5"],
                    help="headline workload: c2 is the configuration the metric is quoted on (default); the others are reported under `workloads` of the default run")
    ap.add_argument("--no-workloads", action="store_true", help="skip the c1/c3/c4/c5 sub-records of the 1-GPU run")
    ap.add_argument("--only-workloads", default=None, help="comma list: restrict the sub-records (profiling runs)")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs: resident steps only (also skips the ordered-rows leg, the sustained loop and the sub-records' extras)")
    ap.add_argument("--e2e-bundles", type=int, default=6, help="bundles per step on the end-to-end path")
    ap.add_argument("--e2e-rows", default="compact8", choices=["compact8", "compact", "packed", "columns"],
                    help="row layout on the end-to-end path: BRB_OPT_PACKED_ROWS = 3 (8-byte rows + 16-bit unit sizes), = 2 (12-byte rows), = 1 (16-byte records) or the plain columns")
    ap.add_argument("--e2e-input", default="compact", choices=["compact", "columns"],
                    help="input columns on the end-to-end path: the compact ones of brb_batch_in (pos0, cigar_cnt16, mate_delta16, refid16) or the plain columns")
    ap.add_argument("--no-bind", action="store_true", help="N>1: do not pin each rank to the CPUs of its GPU's NUMA node")
    ap.add_argument("--e2e-contexts", type=int, default=3, help="brb_ctx (host threads / streams) on the end-to-end path")
    ap.add_argument("--sustained-s", type=float, default=2.0, help="length of the sustained resident loop")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cores = os.cpu_count() or 1
    from bramble_b200 import workloads
    if args.names is None:
        args.names = workloads.DEFAULT_NAMES[args.workload]
    if args.workload in ("c5",) and args.ref_names == 150_000:
        args.ref_names = 30_000
    if args.workload in workloads.LONG and args.ref_names == 150_000:
        args.ref_names = 12_000
    config = {"id": args.workload, "workload": f"{workloads.SHAPE[args.workload]}, {args.names} read names per GPU",
              "flags": " ".join(workloads.FLAGS[args.workload] + (["-S genome.fa"] if args.workload == "c3" else [])) or "(default)", "names_per_gpu": args.names,
              "l2": "inputs + intermediates (>1 GB per step) exceed the 126 MB L2"}

    if args.impl == "reference":
        if rank != 0:
            return
        W = workloads.build(args.workload, 1000, max(args.ref_names, 2000), sample_names=args.ref_names)
        vals, res = [], None
        rb = RefBench(W, {"small": min(2000, args.ref_names // 4), "big": args.ref_names}) if os.path.exists(REF_BIN) else None
        if rb is None:
            print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/bramble_ref not built (reference sources absent)"}))
            return
        pN = max(1, min(cores, 255))
        for _ in range(max(1, min(args.steps, 3))):
            res = rb.leg("big", ["-p", str(pN)])
            if res is None:
                print(json.dumps({"impl": "reference", "unavailable": "the sample is too small to resolve a marginal rate"}))
                return
            vals.append(res["value"])
        v = max(vals)
        print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                          "warmup": args.warmup, "ms_per_step": 1e3 * res["seconds"], "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": v, "unit": UNIT, "cores": pN, "kind": "reference", "sample": rb.sample_text(), "note": SHIM_NOTE},
                          "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        rb.cleanup()
        return

    # stdout carries exactly one JSON line: whatever libraries print there (NCCL's version banner under torchrun) goes to
    # stderr instead; the line itself is written to the saved descriptor at the end
    sys.stdout.flush()
    out_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(line):
        sys.stdout.flush()
        os.write(out_fd, (json.dumps(line) + "\n").encode())

    t_wall0 = time.perf_counter()
    # the long-read bundles are generated on worker processes (fork): before CUDA is initialised in this one
    sub_names = []
    if world == 1 and not args.no_workloads:
        sub_names = [w for w in ("c1", "c3", "c4", "c5") if w != args.workload]
        if args.only_workloads is not None:
            sub_names = [w for w in sub_names if w in args.only_workloads.split(",")]
    with_cpu = not args.no_cpu_baseline and not args.no_e2e
    pre = {w: workloads.build(w, 1000, sample_names=16_000 if with_cpu else 0) for w in sub_names if w in workloads.LONG}
    Wh = workloads.build(args.workload, 1000 + rank, args.names, sample_names=2 * args.ref_names if (with_cpu and rank == 0) else 0)

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
    A, ix, B, p = Wh.A, Wh.ix, Wh.B, Wh.params
    n_aln = int(B.desc.n_aln)
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
    ms, launches, phase = timed_resident(eng, p, args.steps)
    gc.enable()
    barrier()
    clocks = sampler.stop()
    # a sustained loop beside the burst: the same step back to back for >= --sustained-s seconds (power and clocks settle)
    sustained = None
    if not args.no_e2e and args.sustained_s > 0:
        n_sus = max(args.steps, int(args.sustained_s * 1e3 / max(ms / args.steps, 0.05)) + 1)
        s2 = ClockSampler(local)
        s2.prepare()
        gc.disable()
        s2.start()
        ms_sus, _, _ = timed_resident(eng, p, n_sus)
        gc.enable()
        c2 = s2.stop()
        sustained = {"value": n_aln * n_sus / (ms_sus / 1e3), "unit": UNIT, "ms_per_step": ms_sus / n_sus, "steps": n_sus, "seconds": ms_sus / 1e3,
                     "clocks": c2, "what": "rank 0's shard, resident steps back to back for >= %.1f s (the 20-step `value` is a burst)" % args.sustained_s}
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
        if not Wh.long:
            eng.set_option(capi.OPT_PACKED_ROWS, 2)
            eng.run_resident(p)
            checks["compact_rows"] = engine.rows_checksum(eng.download(B, raw=True), batch=B)
            eng.set_option(capi.OPT_PACKED_ROWS, 3)
            eng.run_resident(p)
            checks["compact8_rows"] = engine.rows_checksum(eng.download(B, raw=True), batch=B)
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
        PBs = [eng.pinned_copy(B.slice_groups(g0, g1).compact() if args.e2e_input == "compact" else B.slice_groups(g0, g1)) for g0, g1 in zip(cuts[:-1], cuts[1:])]
        ctxs = [eng.context() for _ in range(max(1, args.e2e_contexts))]
        e2e_packed = 0 if (Wh.long or args.e2e_rows == "columns") else {"compact8": 3, "compact": 2, "packed": 1}[args.e2e_rows]
        for c in ctxs:
            c.set_option(capi.OPT_PACKED_ROWS, e2e_packed)
        eng.set_option(capi.OPT_PACKED_ROWS, e2e_packed)
        engine.project_pipelined(ctxs, p, PBs)
        res = engine.project_pipelined(ctxs, p, PBs, checksum=True)
        checks["e2e_bundles"] = (sum(r[4] for r in res) & 0xFFFFFFFFFFFFFFFF, sum(r[0] for r in res))
        barrier()
        t0 = time.perf_counter()
        # the K steps as ONE stream of K x bundles calls (every call = H2D of its bundle + kernels + D2H of its rows): a host that streams
        # bundles does not drain its pipeline between passes, and neither does the reference's bundle queue
        res = engine.project_pipelined(ctxs, p, PBs * args.steps)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        res = res[:len(PBs)]
        e2e_h2d, e2e_d2h, e2e_rec = sum(r[2] for r in res), sum(r[3] for r in res), sum(r[0] for r in res)
        barrier()
        # the same through ONE call on one context (no overlap), for comparison
        PB = eng.pinned_copy(B.compact() if args.e2e_input == "compact" else B)
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
    line = None
    if rank == 0:
        peak, peak_src = _peaks()
        step_ms = ms_max / args.steps
        value = total_aln * args.steps / (ms_max / 1e3)
        e2e = total_aln * args.steps / (e2e_ms_max / 1e3) if not args.no_e2e else None
        alg = alg_ref
        my_step_ms = ms / args.steps
        achieved = alg / (my_step_ms / 1e3) / 1e9
        names = KERNEL_OF["long" if Wh.long else "short"]
        top = max(phase, key=lambda k: phase[k])
        traffic, traffic_by_kernel, traffic_src = _traffic(args.workload, args.names)
        alg_of = {"ms_lookup": alg_front, "ms_pair": alg - alg_front, "ms_merge": alg - alg_front}
        top_ms = phase[top]
        top_alg = alg_of.get(top, alg) if not Wh.long else alg
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": max(3, args.warmup),
                "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
                "data": "synthetic", "config": config, "gpu_launches": launches, "clocks": clocks,
                "e2e": None if args.no_e2e else {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": e2e_h2d, "d2h_bytes_per_step": e2e_d2h,
                        "records_per_step": e2e_rec, "bundles_per_step": nbun, "contexts": len(ctxs),
                        "how": "brb_project() per bundle from pinned host columns, one host thread per context; the steps follow one another without draining the pipeline",
                        "input": "compact columns (pos0, cigar_cnt16, mate_delta16, refid16)" if args.e2e_input == "compact" else "plain columns",
                        "rows": {3: "BRB_OPT_PACKED_ROWS = 3: 8 bytes per row (tid 22 | pos 24 | flags | CIGAR length), 16-bit unit sizes + one row offset per 32 alignments",
                                 2: "BRB_OPT_PACKED_ROWS = 2: 12 bytes per row (tid, pos, word), 16-bit unit sizes + one row offset per 32 alignments", 1: "BRB_OPT_PACKED_ROWS = 1: one 16-byte record per row",
                                 0: "columns (22 bytes per row + CIGAR CSR)"}[e2e_packed],
                        "single_call_value": n_aln / e2e_single_s, "single_call_note": "rank 0, one brb_project() call for the whole step, no overlap"},
                "sustained": sustained,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "basis": "ALGORITHMIC bytes (SURVEY 8(d): 56 B per reference-defined index hit on every read exon, counted once by the generic kernel) / "
                                      "CUDA-event time of the timed kernel sequence; the timed short-read kernel looks up exon 0 only, so `achieved` is not "
                                      "sustained DRAM bandwidth -- that is `measured_dram_gbs` (ncu DRAM bytes of the same step / the same time)",
                             "traffic": traffic,
                             "measured_dram_gbs": None if traffic is None else traffic / (my_step_ms / 1e3) / 1e9,
                             "measured_dram_frac": None if traffic is None else traffic / (my_step_ms / 1e3) / 1e9 / peak,
                             "traffic_by_kernel": traffic_by_kernel, "traffic_source": f"{traffic_src} (ncu --set full, one step, same kernel sources / workload / size)" if traffic_src else None,
                             "engine_src_sha": engine_src_sha(),
                             "peak_source": peak_src, "scope": "whole kernel sequence of one step (algorithmic bytes of SURVEY 8(d) / DESIGN.md)",
                             "alg_bytes_per_step": alg, "alg_bytes_per_alignment": alg / max(1, n_aln),
                             "top_kernel": names.get(top, top), "top_kernel_share": phase[top] / max(1e-9, sum(phase.values())),
                             "top_kernel_ms": top_ms, "top_kernel_alg_bytes": top_alg,
                             "top_kernel_achieved": top_alg / (top_ms / 1e3) / 1e9, "top_kernel_frac": top_alg / (top_ms / 1e3) / 1e9 / peak,
                             "phase_ms_per_step": {names[k]: v for k, v in phase.items() if k in names},
                             "ordered_rows": None if ms_ordered is None else {"ms_per_step": ms_ordered / args.steps, "frac": alg / (ms_ordered / args.steps / 1e3) / 1e9 / peak,
                                              "what": "rank 0, BRB_OPT_ORDERED_ROWS=1: rows moved into emit order on the device (scan + k_place + k_group_rows) "
                                                      "instead of unit blocks + aln_rec_off/aln_rec_cnt"},
                             "packed_rows": None if ms_packed is None else {"ms_per_step": ms_packed / args.steps, "frac": alg / (ms_packed / args.steps / 1e3) / 1e9 / peak,
                                              "what": "rank 0, BRB_OPT_PACKED_ROWS=1: the same rows written as 16-byte records (same algorithmic bytes: they are "
                                                      "defined on the reference's row content, not on this layout)"}},
                "parity_at_bench_size": None if not checks else {
                    "property": "order-independent checksum of all rows (source alignment, tid, pos, strand, flags, scores, CIGAR) and row count agree between "
                                "the default kernels, the generic front kernel, BRB_OPT_ORDERED_ROWS, BRB_OPT_PACKED_ROWS and the 8-bundle / 3-context end-to-end path (rank 0); "
                                "`oracle` = the same checksum from the CPU oracle on the first 200k read names",
                    "rows": checks["default"][1], "checksum": "%016x" % checks["default"][0],
                    "agree": len({(int(c[0]), int(c[1])) for c in checks.values()}) == 1},
                "host": {**host, "e2e_ms_per_step_by_rank": [round(float(x[0].item()), 3) for x in by_rank],
                         "numa_node_by_rank": [int(x[1].item()) for x in by_rank], "cpus_allowed_by_rank": [int(x[2].item()) for x in by_rank],
                         "ms_per_step_by_rank": [round(float(x[3].item()), 4) for x in by_rank], "sm_mhz_by_rank": [int(x[4].item()) for x in by_rank],
                         "slowdown_seen_by_rank": [bool(x[5].item()) for x in by_rank]},
                "work": {"alignments_per_gpu": n_aln, "candidates": int(st.n_candidates), "matches": int(st.n_matches), "records": int(st.n_rec)}}
        if not args.no_e2e:
            try:
                op = oracle_parity(eng, Wh)
                line["parity_at_bench_size"]["oracle"] = op
                line["parity_at_bench_size"]["oracle_agree"] = op["oracle_agree"]
            except Exception as e:
                line["parity_at_bench_size"]["oracle_error"] = str(e)[:300]
    # the headline engine is done: release its device memory before the other workloads
    if PB is not None:
        eng.pinned_free(PB)
    for x in PBs:
        eng.pinned_free(x)
    for c in ctxs:
        c.close()
    eng.close()
    if rank == 0:
        rb = None
        if with_cpu and os.path.exists(REF_BIN):
            try:
                G = int(B.desc.n_groups)
                samples = {"small": min(2000, args.ref_names // 4), "cli": min(2 * args.ref_names, G)}     # the CLI gets a larger sample: it needs several bundles in flight
                if args.gpus == 1:
                    samples.update(big=min(args.ref_names, G), mid=min(max(args.ref_names // 3, 4000), G))
                rb = RefBench(Wh, samples)
            except Exception as e:
                line["cpu_baseline_error"] = str(e)[:300]
        if rb is not None and args.gpus == 1:
            try:
                cb = cpu_baseline(rb, cores, full_legs=True)
                if cb:
                    if line["parity_at_bench_size"] and line["parity_at_bench_size"].get("oracle"):
                        cb["port_projection_only"] = line["parity_at_bench_size"]["oracle"]["oracle_port"]
                    line["cpu_baseline"] = cb
            except Exception as e:     # the baseline is a reported number, never a reason to lose the bench line
                line["cpu_baseline_error"] = str(e)[:300]
        if rb is not None:
            # the drop-in CLI end to end INCLUDING BGZF, at 1 GPU and -- when this run spans several -- at every power of two up
            # to N (one process, the CLI's own dealer, driven from rank 0 while the other ranks wait at the barrier below)
            try:
                gl = [n for n in (1, 2, 4, 8) if n <= max(1, args.gpus)]
                ce = cli_e2e(rb, cores, gl, repeats=3 if args.gpus == 1 else 2)
                if "1" in ce:
                    line["cli_e2e"] = dict(ce["1"], what="bramble_b200_cli BAM in -> BAM out (1 GPU): alignments of the sample / the CLI's pipeline time "
                                           "(BRB_CLI_TIMING: first input block to closed output), best of the runs; marginal_value = the same by the reference "
                                           "arm's method (difference of two process wall times)")
                if len(ce) > 1:
                    line["cli_e2e_by_gpus"] = ce
            except Exception as e:
                line["cli_e2e_error"] = str(e)[:300]
            rb.cleanup()
        if sub_names:
            line["workloads"] = {}
            for w in sub_names:
                try:
                    Ws = pre.pop(w) if w in pre else workloads.build(w, 1000)
                    line["workloads"][w] = measure_workload(Ws, args, local, cores, peak, with_cpu=with_cpu)
                    del Ws
                except Exception as e:
                    line["workloads"][w] = {"error": str(e)[:300]}
            if line.get("parity_at_bench_size") is not None:
                line["parity_at_bench_size"]["oracle_agree_all_workloads"] = bool(line["parity_at_bench_size"].get("oracle_agree")) and all(
                    v.get("parity_at_bench_size", {}).get("oracle_agree") is True for v in line["workloads"].values())
        line["bench_wall_s"] = time.perf_counter() - t_wall0
        emit(line)
    barrier()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
