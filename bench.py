#!/usr/bin/env python
"""bench.py -- Spartan NIZK prove throughput (constraints/s) on B200, per the driver contract.

A "step" is one NIZK::prove of one synthetic R1CS (Instance::produce_synthetic_r1cs, r1csinstance.rs:136-214; 10 inputs,
one non-zero per row and matrix, uniform 253-bit witness -- the worst case for the commitment MSM).  The default workload
is the one north_star's Target names: 2^22 constraints on one GPU (configs[4]'s single-GPU size); --log-size 16 gives
configs[1].  Generators and instance are built outside the timed region like benches/nizk.rs:27-39, and reported beside
it under "cold", because the reference's own caller (spartan-zkinterface/src/main.rs:71-85) times NIZKGens::new with the
prove.

  value   witness already resident in HBM when the timed region starts (CUDA events on the library's stream)
  e2e     the same prove through the public API (otti_b200.NIZK.prove) from PINNED host memory: host->device copy of the
          witness and device->host read of the proof inside the timed region.  Resident and e2e steps alternate (ABAB),
          each with its own L2 flush; median / p10 / p90 of both arms are reported.
  cold    one-shot generators (sb200_gens_new_for(..., proofs_expected = 1)) + the first prove + verify, wall clock
  side    (N = 1) configs[1] (2^16) and an Otti-shaped 2^20 instance with a heavy constant column through
          zkinterface ingestion -> Instance::new -> one-shot generators -> prove -> verify
  N > 1   one process per GPU, each proving its own instance (independent proofs are the data-parallel unit of Otti's
          run.py loop over datasets: no data-path collective, weak scaling) + ONE proof sharded over all N GPUs.

--impl reference times the CPU restatement of the reference prover (oracle/, "port": the real curve25519-dalek prover is
Rust and cannot be built in this image) with all host threads, on a bounded sample of the same workload.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

LABEL = b"nizk_example"
TAPE = bytes(range(64))
SEED = 0x0771
NUM_INPUTS = 10
README_2_20_CONSTRAINTS_PER_S = (1 << 20) / 4.5140  # Spartan/README.md:365-377 (1 core, i7-1065G7)
METRIC = "spartan_nizk_prove_constraints_per_s"
DTYPE = "u256 (F_l 253-bit / F_p 255-bit integers)"


def workload_name(s):
    return "synthetic Spartan R1CS 2^%d constraints = vars, 10 inputs, NIZK::prove" % s


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log-size", type=int, default=22, help="log2 of constraints = vars (north_star Target = 22, configs[1] = 16)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-side", action="store_true", help="skip the side entries (2^16, Otti-shaped 2^20, SNARK mode; N = 1 only)")
    ap.add_argument("--no-snark", action="store_true", help="skip the SNARK-mode side measurements")
    ap.add_argument("--otti-log-size", type=int, default=20, help="log2 constraints of the Otti-shaped side entry")
    ap.add_argument("--shard-log-sizes", default="22,24",
                    help="N > 1: log2 sizes of the ONE proof additionally sharded over all N GPUs (configs[4]); empty = skip")
    ap.add_argument("--exchange", default="shm", choices=["shm", "nccl", "nccl-native"],
                    help="exchange hook of the sharded proof: POSIX shared memory, torch.distributed (NCCL through a Python callback), ncclAllGather called natively")
    return ap.parse_args()


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def quantiles(xs):
    ys = sorted(xs)
    n = len(ys)
    q = lambda f: ys[min(n - 1, max(0, int(round(f * (n - 1)))))]
    return {"median": q(0.5), "p10": q(0.1), "p90": q(0.9), "min": ys[0], "max": ys[-1], "n": n}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md:70-72).  The sampler process is
    moved to the LAST cpu of the box, outside every rank's slice (sb200_init binds ranks to the cores next to their GPU)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.p = gpu, None

    def start(self):
        try:
            def unbind():
                try:
                    ncpu = os.cpu_count() or 1
                    os.sched_setaffinity(0, {ncpu - 1})
                    os.nice(10)
                except Exception:
                    pass
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, preexec_fn=unbind)
        except Exception:
            self.p = None

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            out = self.p.communicate(timeout=5)[0]
        except Exception:
            self.p.kill()
            out = ""
        sm, mx, reasons = [], [], set()
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    except Exception:
        return 6650.0, "fallback of B200_PROFILING.md (no MEASURED_PEAKS.json)"


_OUT_FD = None


def claim_stdout():
    """Keep the process's stdout for the ONE JSON line: everything else that writes to file descriptor 1 (NCCL prints its
    version banner there from C, whatever NCCL_DEBUG_FILE says) goes to stderr from here on."""
    global _OUT_FD
    if _OUT_FD is None:
        sys.stdout.flush()
        _OUT_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _OUT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_OUT_FD, data)


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args, rank, world):
    """CPU restatement of the reference prover on the host cores (rank 0 only): each step is one full prove of a bounded
    sample of the workload (2^min(s, 20): the prover is O(N), constraints/s is size-independent to first order)."""
    if rank != 0:
        return
    import oracle_lib as O
    s = min(args.log_size, 20)
    n = 1 << s
    threads = host_threads()
    inst, vars_, inputs = O.Instance.synthetic(n, n, NUM_INPUTS, SEED)
    gens = O.Gens(n, n, NUM_INPUTS)
    for _ in range(min(args.warmup, 1)):
        O.nizk_prove(inst, vars_, inputs, gens, LABEL, TAPE, threads)
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        O.nizk_prove(inst, vars_, inputs, gens, LABEL, TAPE, threads)
        times.append(time.perf_counter() - t0)
    dt = sum(times)
    value = n * args.steps / dt
    sample = "%d full proves of synthetic R1CS 2^%d on %d host threads (commitment rows threaded like the reference's `multicore` feature)" % (
        args.steps, s, threads)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "constraints/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
        "config": {"workload": workload_name(args.log_size), "sample_log_size": s,
                   "note": "CPU restatement (oracle/, kind 'port'), not curve25519-dalek: the Rust toolchain is absent"},
        "step_ms": quantiles([1e3 * t for t in times]),
        "cpu_baseline": {"value": value, "unit": "constraints/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "constraints/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ------------------------------------------------------------------------------------------------ helpers (our arm)
class Env:
    """torch / library handles shared by the measurement functions"""

    def __init__(self, args, torch, sb, dist, rank, world, local_rank):
        self.args, self.torch, self.sb, self.dist = args, torch, sb, dist
        self.rank, self.world, self.local_rank = rank, world, local_rank
        self.flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def flush_l2(self, i):
        self.flush_buf.fill_(i & 0xFF)
        self.torch.cuda.synchronize()

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()


def timed_abab(env, inst, vars_, inputs, gens, steps, comm=None):
    """K resident steps and K end-to-end steps, alternating; each bracketed by CUDA events on the library's stream."""
    torch, sb = env.torch, env.sb
    ext = torch.cuda.ExternalStream(inst.stream)
    res, e2e, launches, profs, profs_e2e, proof = [], [], 0, [], [], None
    for i in range(steps):
        for resident in (True, False):
            env.flush_l2(2 * i + (0 if resident else 1))
            if comm is not None:
                env.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ext)
            proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE, resident=resident, comm=comm)
            e1.record(ext)
            e1.synchronize()
            ms = e0.elapsed_time(e1)
            prof = sb.NIZK.last_profile()
            if resident:
                res.append(ms)
                profs.append(prof)
                launches += int(prof["kernel_launches"])
            else:
                e2e.append(ms)
                profs_e2e.append(prof)
    return res, e2e, launches, profs, profs_e2e, proof


def kernel_profile(sb, run, reps=2):
    """per-kernel device time of `reps` calls of run() (CUDA events around every launch, outside any timed region):
    [(kernel, launches per call, ms per call)] longest first"""
    L = sb.lib()
    L.sb200_k_kprof_enable(1)
    buf = C.create_string_buffer(1 << 16)
    L.sb200_k_kprof_report(buf, C.c_size_t(len(buf)))  # drop whatever was pending
    for _ in range(reps):
        run()
    L.sb200_k_kprof_report(buf, C.c_size_t(len(buf)))
    L.sb200_k_kprof_enable(0)
    out = []
    for line in buf.value.decode().splitlines():
        name, cnt, ms = line.rsplit(",", 2)
        out.append((name, int(cnt) / reps, float(ms) / reps))
    return out


def int_pipe_peak(sm_mhz):
    """IMAD.WIDE.U32 issue rate of the whole chip.  One warp instruction per 4 cycles per SM sub-partition = 8 lanes per
    clock per sub-partition: measured with tools/microbench/fpmul_bench under ncu (profiles/r02a_pipe_fpmul.csv + r02a_pipe_msm_fixed_2p22.csv)."""
    return 148 * 4 * 8 * sm_mhz * 1e6


def roofline_for(kernel, kernel_s, share, s, c, sm_mhz, extra=None):
    """Roofline record of the dominant kernel.  Algorithmic work per launch:
      msm_fixed_kernel   (N_v + L) terms x W windows x (1 - 2^-c) mixed additions x 7 F_p products x 75 IMAD.WIDE.U32;
                         bytes = 32 per term + 96 per mixed addition (one table entry)
      msm_small_kernel   (R + 2) terms x W windows x rows, same per-addition figures (latency-bound tree reduction)
      sc_*_kernel        4 (cubic) / 2 (quadratic) tables x 32 B x (read n + write n/2) per round"""
    n = 1 << s
    ell = s
    Ls, Rs = 1 << (ell // 2), 1 << (ell - ell // 2)
    W = 252 // c + 1
    hbm_peak, which = measured_peaks()
    rec = {"kernel": kernel, "kernel_ms": 1e3 * kernel_s, "share_of_step": share}
    if kernel.startswith("msm_fixed") or kernel.startswith("msm_small"):
        if kernel.startswith("msm_fixed"):
            terms, rows_note = n + Ls, "commitment rows %d x %d" % (Ls, Rs)
        else:
            terms, rows_note = (extra or {}).get("terms", Rs + 2), "small MSM rows over %d bases" % (Rs + 2)
        madds = terms * W * (1 - 2.0 ** -c)
        alg_bytes = terms * 32 + madds * 96
        imad = madds * 7 * 75
        peak = int_pipe_peak(sm_mhz)
        rec.update({
            "kernel": "%s (%s, %d-bit windows)" % (kernel, rows_note, c), "bound": "int_pipe",
            "achieved": imad / kernel_s / 1e9 if kernel_s > 0 else 0.0, "peak": peak / 1e9, "unit": "GIMAD/s (IMAD.WIDE.U32 lane-operations)",
            "frac": (imad / kernel_s) / peak if kernel_s > 0 else 0.0,
            "peak_source": "148 SM x 4 sub-partitions x 8 lanes/clk x %.0f MHz (sampled under load); 1 IMAD.WIDE.U32 warp instruction "
                           "per 4 cycles per sub-partition, profiles/r02a_pipe_fpmul.csv + r02a_pipe_msm_fixed_2p22.csv" % sm_mhz,
            "algorithmic_imad": imad, "algorithmic_bytes": alg_bytes,
            "hbm": {"bound": "hbm", "achieved": alg_bytes / kernel_s / 1e9 if kernel_s > 0 else 0.0, "peak": hbm_peak, "unit": "GB/s",
                    "frac": alg_bytes / kernel_s / 1e9 / hbm_peak if kernel_s > 0 else 0.0, "peak_source": which},
        })
    else:
        ntab = 3 if "cubic_eq" in kernel else 4 if "cubic" in kernel else 2
        rec.update({"bound": "hbm", "unit": "GB/s", "peak": hbm_peak, "peak_source": which, "achieved": None, "frac": None,
                    "note": "per-round bytes = %d tables x 32 B x 1.5 n; see hbm_kernel_sample for round 0" % ntab})
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")) as f:
            ent = json.load(f).get(kernel.split(" ")[0], {}).get(str(s))
        if ent:
            traffic, traffic_src = ent["bytes"], ent["source"]
    except Exception:
        pass
    rec["traffic"] = traffic
    rec["traffic_unit"] = "bytes per launch (ncu dram__bytes_read.sum + dram__bytes_write.sum)"
    rec["traffic_source"] = traffic_src
    return rec


def cold_path(env, n, inst, vars_, inputs, t_inst):
    """What a one-shot caller pays (spzk: the timer starts before NIZKGens::new, main.rs:71-85, and verify is printed
    right after, :89-103): generators sized for ONE proof, the first prove on a fresh workspace, verify.  Wall clock."""
    sb, torch = env.sb, env.torch
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    g1 = sb.NIZKGens.new(n, n, NUM_INPUTS, proofs_expected=1)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    proof = sb.NIZK.prove(inst, vars_, inputs, g1, sb.Transcript.new(LABEL), TAPE)
    t2 = time.perf_counter()
    proof2 = sb.NIZK.prove(inst, vars_, inputs, g1, sb.Transcript.new(LABEL), TAPE)
    t3 = time.perf_counter()
    proof.verify(inst, inputs, sb.Transcript.new(LABEL), g1)
    t4 = time.perf_counter()
    out = {"instance_synthetic_s": t_inst, "gens_build_s": t1 - t0, "gens_window_bits": g1.window_bits,
           "first_prove_s": t2 - t1, "second_prove_s": t3 - t2, "verify_s": t4 - t3,
           "gens_plus_prove_s": (t1 - t0) + (t2 - t1),
           "note": "generators built with proofs_expected = 1 (sb200_gens_new_for); first_prove_s includes the one-off "
                   "workspace allocation of the instance; instance_synthetic_s = R1CS generation + Instance::new",
           "proof_same_as_steady_state": None}
    return out, proof2.to_bytes(), g1


def side_2p16(env):
    """configs[1]: synthetic 2^16, steady state (ABAB, 30 steps) + the cold path"""
    sb, torch = env.sb, env.torch
    s, n = 16, 1 << 16
    t0 = time.perf_counter()
    inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, NUM_INPUTS, SEED, pinned=True)
    t_inst = time.perf_counter() - t0
    cold, cold_bytes, g1 = cold_path(env, n, inst, vars_, inputs, t_inst)
    del g1
    t0 = time.perf_counter()
    gens = sb.NIZKGens.new(n, n, NUM_INPUTS)
    torch.cuda.synchronize()
    gens_s = time.perf_counter() - t0
    for _ in range(3):
        sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE)
    inst.upload_witness(gens, vars_)
    res, e2e, launches, profs, _, proof = timed_abab(env, inst, vars_, inputs, gens, 30)
    cold["proof_same_as_steady_state"] = cold_bytes == proof.to_bytes()
    kp = kernel_profile(sb, lambda: sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE, resident=True), 3)
    step_ms = sum(res) / len(res)
    out = {"workload": workload_name(s) + " (configs[1])", "steps": 30, "prove_ms": quantiles(res), "e2e_prove_ms": quantiles(e2e),
           "ms_per_step": step_ms, "constraints_per_s": n / (step_ms * 1e-3), "launches_per_proof": launches / 30.0,
           "gens_build_s_many_proofs": gens_s, "msm_window_bits": gens.window_bits, "cold": cold,
           "kernels_ms_per_proof": [{"kernel": k, "launches": c, "ms": round(ms, 4)} for k, c, ms in kp[:8]],
           "kernel_time_share_of_step": sum(ms for _, _, ms in kp) / step_ms}
    if kp:
        k, cnt, ms = kp[0]
        Rs = 1 << 8
        out["roofline"] = roofline_for(k, ms * 1e-3 / max(cnt, 1), ms / step_ms, s, gens.window_bits, 1965.0,
                                       {"terms": 2 * (Rs + 2)} if k.startswith("msm_small") else None)
        out["roofline"]["note"] = ("per launch; the step is latency-bound at this size (33 sum-check + 8 bullet Fiat-Shamir rounds), "
                                   "no kernel is near a device roofline")
    snark = None
    if not env.args.no_snark:
        snark = snark_side_measurements(sb, inst, n, s)
    return out, snark


def side_otti(env, s):
    """Otti-shaped instance (tools/zkifgen.cpp: 80 % bits, 30 % of all entries in the constant column) of 2^s constraints
    through the reference caller's whole sequence: zkinterface parse -> Instance::new -> NIZKGens::new -> prove -> verify."""
    sb, torch = env.sb, env.torch
    import zkifgen
    from otti_b200 import zkif
    n = 1 << s
    g = zkifgen.OttiShaped(n, n, NUM_INPUTS, seed=11)
    t0 = time.perf_counter()
    r = zkif.R1cs.from_pointers(g.file_ptr(0), g.file_ptr(1), g.file_ptr(2))
    t1 = time.perf_counter()
    inst = r.instance()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    vars_, inputs = r.vars_assignment(), r.inputs_assignment()
    t2b = time.perf_counter()
    gens = r.nizk_public_params(proofs_expected=1)
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE)
    t4 = time.perf_counter()
    proof.verify(inst, inputs, sb.Transcript.new(LABEL), gens)
    t5 = time.perf_counter()
    steady, vsteady = [], []
    for _ in range(5):
        ta = time.perf_counter()
        p2 = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE)
        steady.append(1e3 * (time.perf_counter() - ta))
        ta = time.perf_counter()
        p2.verify(inst, inputs, sb.Transcript.new(LABEL), gens)
        vsteady.append(1e3 * (time.perf_counter() - ta))
    kp = kernel_profile(sb, lambda: sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE), 2)
    kd = {k: (c, ms) for k, c, ms in kp}
    nnz = sum(g.nnz)
    hbm_peak, _ = measured_peaks()
    sp = {}
    for name, bytes_ in (("spmv3_kernel", 36 * nnz + 3 * 4 * (n + 1) + 3 * 32 * n + 64 * n),
                         ("spmvT3_kernel", 36 * nnz + 3 * 4 * (2 * n + 1) + 32 * n + 64 * n)):
        longk = sum(kd.get(k, (0, 0.0))[1] for k in ("sp_long_kernel", "sp_big_kernel")) / 2.0
        ms = kd.get(name, (0, 0.0))[1] + longk
        sp[name] = {"ms_with_long_segment_passes": round(ms, 4), "algorithmic_bytes": bytes_,
                    "gb_per_s": bytes_ / (ms * 1e-3) / 1e9 if ms > 0 else None,
                    "frac_of_hbm_peak": bytes_ / (ms * 1e-3) / 1e9 / hbm_peak if ms > 0 else None}
    out = {"workload": "Otti-shaped R1CS 2^%d constraints = vars (shape-substituted configs[3]; tools/zkifgen.cpp), through zkinterface" % s,
           "nnz": list(g.nnz), "constant_column_share_of_nnz": g.const_entries / float(nnz),
           "zkif_bytes": g.file_ptr(0)[1] + g.file_ptr(1)[1] + g.file_ptr(2)[1],
           "cold": {"zkif_parse_s": t1 - t0, "instance_new_s": t2 - t1, "assignments_s": t2b - t2, "gens_build_s": t3 - t2b,
                    "gens_window_bits": gens.window_bits, "prove_s": t4 - t3, "verify_s": t5 - t4,
                    "gens_plus_prove_s": (t3 - t2b) + (t4 - t3), "total_s": t5 - t0},
           "steady_prove_ms": quantiles(steady), "steady_verify_ms": quantiles(vsteady),
           "constraints_per_s_steady": n / (quantiles(steady)["median"] * 1e-3),
           "proof_bytes": len(proof.to_bytes()), "proof_deterministic": p2.to_bytes() == proof.to_bytes(),
           "sparse_kernels": sp,
           "kernels_ms_per_proof": [{"kernel": k, "launches": c, "ms": round(ms, 4)} for k, c, ms in kp[:8]]}
    if not env.args.no_cpu_baseline:
        import oracle_lib as O
        oi = O.Instance.new_arrays(n, n, NUM_INPUTS, [g.coo_arrays(m) + (g.nnz[m],) for m in range(3)])
        og = O.Gens(n, n, NUM_INPUTS)
        threads = host_threads()
        ta = time.perf_counter()
        want = O.nizk_prove(oi, g.vars_bytes(), g.inputs_bytes(), og, LABEL, TAPE, threads)
        tb = time.perf_counter()
        ok = O.nizk_verify(oi, g.inputs_bytes(), og, LABEL, proof.to_bytes()) == 0
        tc = time.perf_counter()
        out["cpu_port"] = {"prove_s": tb - ta, "verify_s": tc - tb, "cores": threads, "proof_bytes_equal_gpu": want == proof.to_bytes(),
                           "oracle_verify_accepts_gpu_proof": ok}
    return out


def side_otti_sweep(env, sizes=(12, 14, 16, 18)):
    """The Otti suites span 2^12 .. 2^20 constraints (SURVEY.md section 8(d) ii): the same Otti-shaped generator at the smaller
    sizes, compact records -- one-shot path (parse, Instance::new, generators for one proof, prove, verify) and steady prove /
    verify -- with the CPU restatement's prove beside them."""
    sb, torch = env.sb, env.torch
    import statistics
    import zkifgen
    from otti_b200 import zkif
    out = []
    for s in sizes:
        n = 1 << s
        g = zkifgen.OttiShaped(n, n, NUM_INPUTS, seed=11)
        t0 = time.perf_counter()
        r = zkif.R1cs.from_pointers(g.file_ptr(0), g.file_ptr(1), g.file_ptr(2))
        inst = r.instance()
        vars_, inputs = r.vars_assignment(), r.inputs_assignment()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        gens = r.nizk_public_params(proofs_expected=1)
        proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE)
        t2 = time.perf_counter()
        proof.verify(inst, inputs, sb.Transcript.new(LABEL), gens)
        t3 = time.perf_counter()
        del gens
        gens = r.nizk_public_params()
        for _ in range(2):
            sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE)
        tp, tv = [], []
        for _ in range(7):
            ta = time.perf_counter()
            p2 = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE)
            tp.append(1e3 * (time.perf_counter() - ta))
            ta = time.perf_counter()
            p2.verify(inst, inputs, sb.Transcript.new(LABEL), gens)
            tv.append(1e3 * (time.perf_counter() - ta))
        rec = {"log_size": s, "nnz": sum(g.nnz), "one_shot": {"ingest_s": round(t1 - t0, 4), "gens_plus_prove_s": round(t2 - t1, 4),
                                                               "verify_s": round(t3 - t2, 4), "total_s": round(t3 - t0, 4)},
               "steady_prove_ms": round(statistics.median(tp), 3), "steady_verify_ms": round(statistics.median(tv), 3),
               "proof_bytes": len(proof.to_bytes()), "proof_same_with_both_generator_tables": p2.to_bytes() == proof.to_bytes()}
        if not env.args.no_cpu_baseline:
            import oracle_lib as O
            oi = O.Instance.new_arrays(n, n, NUM_INPUTS, [g.coo_arrays(m) + (g.nnz[m],) for m in range(3)])
            og = O.Gens(n, n, NUM_INPUTS)
            threads = host_threads()
            ta = time.perf_counter()
            want = O.nizk_prove(oi, g.vars_bytes(), g.inputs_bytes(), og, LABEL, TAPE, threads)
            rec["cpu_port_prove_s"] = round(time.perf_counter() - ta, 4)
            rec["cpu_port_cores"] = threads
            rec["proof_bytes_equal_cpu_port"] = want == proof.to_bytes()
        out.append(rec)
        del gens, inst
    return out


def snark_side_measurements(sb, inst, n, s):
    """SNARK mode on a 2^16 instance, outside the timed region of the headline metric: whole-call
    milliseconds (best of 4) of each step, then SNARK::prove / verify.  Never fails the bench line."""
    try:
        import random
        rng = random.Random(s)
        ell = 2**252 + 27742317777372353535851937790883648493
        sc = lambda k: b"".join(rng.randrange(ell).to_bytes(32, "little") for _ in range(k))
        t0 = time.perf_counter()
        gens = sb.SNARKGens.new(n, n, NUM_INPUTS, n)
        out = {"workload": "synthetic R1CS 2^%d, instance resident" % s,
               "gens_build_s_untimed": round(time.perf_counter() - t0, 3)}
        rx, ry, rh, rm = sc(s), sc(s + 1), sc(1), sc(1)
        calls = {"encode_ms": lambda: sb.SNARK.encode(inst, gens),
                 "commit_nondet_witness_ms": lambda: sb.SNARK.commit_derefs(inst, gens, rx, ry),
                 "build_layered_network_ms": lambda: sb.SNARK.network_products(inst, rx, ry, rh, rm)}
        for key, fn in calls.items():
            best = None
            for _ in range(4):
                fn()
                p = sb.SNARK.last_encode_profile()
                best = p if best is None or p["total_ms"] < best["total_ms"] else best
            out[key] = round(best["total_ms"], 3)
            out[key.replace("_ms", "_device_ms")] = round(best["dense_ms"] + best["commit_ops_ms"] + best["commit_mem_ms"], 3)
        if hasattr(sb.SNARK, "prove"):
            out.update(snark_prove_measurements(sb, inst, gens, n, s))
        out["published_reference"] = {"encode_s": 14.26, "commit_nondet_witness_s": 14.45, "build_layered_network_s": 3.44,
                                      "prove_s": 39.13, "note": "Spartan/README.md:305-345, 2^20, 1 core (other hardware)"}
        return out
    except Exception as e:  # a side measurement must not cost the headline line
        return {"error": "%s: %s" % (type(e).__name__, e)}


SNARK_LABEL = b"snark_example"


def snark_prove_measurements(sb, inst, gens, n, s, vars_=None, inputs=None, oracle_check=True):
    """SNARK::prove / SNARK::verify (lib.rs:335-457) on the resident instance: wall-clock milliseconds of the calls (host
    buffers in, proof bytes out), the scopes of the reference's profiler, the proof against the CPU restatement."""
    import statistics
    if vars_ is None:
        _, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, NUM_INPUTS, SEED)
    comm = sb.SNARK.encode(inst, gens)
    proof = None
    for _ in range(2):
        proof = sb.SNARK.prove(inst, comm, None, vars_, inputs, gens, sb.Transcript.new(SNARK_LABEL), TAPE)
    ts, profs = [], []
    for _ in range(5):
        t0 = time.perf_counter()
        proof = sb.SNARK.prove(inst, comm, None, vars_, inputs, gens, sb.Transcript.new(SNARK_LABEL), TAPE)
        ts.append(1e3 * (time.perf_counter() - t0))
        profs.append(sb.SNARK.last_prove_profile())
    tv = []
    for _ in range(3):
        t0 = time.perf_counter()
        proof.verify(comm, inputs, sb.Transcript.new(SNARK_LABEL), gens)
        tv.append(1e3 * (time.perf_counter() - t0))
    best = profs[ts.index(min(ts))]
    out = {"prove_ms": round(statistics.median(ts), 3), "prove_ms_min": round(min(ts), 3), "verify_ms": round(statistics.median(tv), 3),
           "proof_bytes": len(proof.to_bytes()), "kernel_launches_per_proof": int(best["kernel_launches"]),
           "prove_scopes_ms": {k: round(1e3 * best[k], 3) for k in ("r1cs_sat_proof", "eval_sparse_polys", "commit_nondet_witness",
                                                                    "build_layered_network", "product_layer_proof", "hash_layer_proof")}}
    if oracle_check:
        import oracle_lib as O
        oi, ov, oin = O.Instance.synthetic(n, n, NUM_INPUTS, SEED)
        threads = host_threads()
        t0 = time.perf_counter()
        want = O.snark_prove(oi, n, ov, oin, SNARK_LABEL, TAPE, threads=threads)
        t1 = time.perf_counter()
        ok = O.snark_verify(comm.to_bytes(), n, oin, SNARK_LABEL, proof.to_bytes()) == 0
        t2 = time.perf_counter()
        out["cpu_port"] = {"prove_s": round(t1 - t0, 3), "verify_s": round(t2 - t1, 3), "cores": threads,
                           "proof_bytes_equal_gpu": want == proof.to_bytes(), "oracle_verify_accepts_gpu_proof": ok}
    return out


def side_snark_2p20(env):
    """SNARK mode at the size the reference publishes its profile for (README.md:305-345: 2^20 constraints, variables and
    non-zero entries per matrix): encode, prove, verify through the public API."""
    sb, torch = env.sb, env.torch
    s, n = 20, 1 << 20
    inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, NUM_INPUTS, SEED, pinned=True)
    t0 = time.perf_counter()
    gens = sb.SNARKGens.new(n, n, NUM_INPUTS, n)
    torch.cuda.synchronize()
    out = {"workload": "synthetic R1CS 2^20 (constraints = variables = non-zero entries per matrix), SNARK mode",
           "gens_build_s_untimed": round(time.perf_counter() - t0, 3)}
    best = None
    for _ in range(3):
        sb.SNARK.encode(inst, gens)
        p = sb.SNARK.last_encode_profile()
        best = p if best is None or p["total_ms"] < best["total_ms"] else best
    out["encode_ms"] = round(best["total_ms"], 3)
    out.update(snark_prove_measurements(sb, inst, gens, n, s, vars_, inputs, oracle_check=not env.args.no_cpu_baseline))
    out["published_reference"] = {"encode_s": 14.26, "prove_s": 39.13, "r1cs_sat_proof_s": 4.44, "eval_sparse_polys_s": 0.377,
                                  "commit_nondet_witness_s": 14.45, "build_layered_network_s": 3.44, "evalproof_layered_network_s": 15.57,
                                  "proof_bytes": 47024 + 96 + 133720,
                                  "note": "Spartan/README.md:305-345, 1 core i7-1065G7 (other hardware; not the same measurement)"}
    return out


def run_sharded(env, s):
    """ONE proof of 2^s constraints sharded over all ranks: constraint rows / z columns cyclic,
    commitment rows in blocks, per-round exchange of 96 bytes per rank; every rank gets the same proof bytes.
    Time = max over ranks of the CUDA-event time around the prove.  Rank 0 then proves the same R1CS alone."""
    import hashlib
    import uuid
    from otti_b200 import comm as sbcomm
    args, sb, torch, dist, rank, world = env.args, env.sb, env.torch, env.dist, env.rank, env.world
    n = 1 << s
    name = ["/sb200-bench-" + uuid.uuid4().hex[:12]]
    dist.broadcast_object_list(name, src=0)
    comm = (sbcomm.ShmComm(name[0], rank, world) if args.exchange == "shm" else
            sbcomm.NcclNativeComm() if args.exchange == "nccl-native" else sbcomm.DistComm())
    inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, NUM_INPUTS, SEED, pinned=True, rank=rank, world=world)
    gens = sb.NIZKGens.new(n, n, NUM_INPUTS)
    proof = None
    for _ in range(args.warmup):
        proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE, comm=comm)
    inst.upload_witness(gens, vars_)
    steps = max(3, min(args.steps, 10))
    res, e2e, _, profs, _, proof = timed_abab(env, inst, vars_, inputs, gens, steps, comm=comm)
    prof = profs[-1]
    digest = hashlib.sha256(proof.to_bytes()).digest()
    t = torch.tensor([sum(res), sum(e2e)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    digests = [None] * world
    dist.all_gather_object(digests, digest)
    comm.close()
    del inst
    single_ms, same_as_single = None, None
    if rank == 0:  # the same R1CS on one GPU, for the speed-up
        inst1, v1, i1 = sb.Instance.produce_synthetic_r1cs(n, n, NUM_INPUTS, SEED, pinned=True)
        for _ in range(args.warmup):
            p1 = sb.NIZK.prove(inst1, v1, i1, gens, sb.Transcript.new(LABEL), TAPE)
        inst1.upload_witness(gens, v1)
        r1, _, _, _, _, p1 = timed_abab(env, inst1, v1, i1, gens, steps)
        single_ms = sum(r1) / len(r1)
        same_as_single = hashlib.sha256(p1.to_bytes()).digest() == digest
        del inst1
    del gens
    env.barrier()
    if rank != 0:
        return None
    ms = float(t[0]) / steps
    return {
        "workload": "ONE synthetic Spartan R1CS 2^%d sharded over %d GPUs (rows/columns cyclic, commitment rows in blocks)" % (s, world),
        "scaling": "strong", "steps": steps, "prove_ms": ms, "constraints_per_s": n / (ms * 1e-3),
        "e2e_prove_ms": float(t[1]) / steps, "single_gpu_prove_ms": single_ms,
        "speedup_vs_single_gpu": (single_ms / ms) if single_ms else None,
        "exchange": args.exchange, "exchange_calls_per_proof": int(prof["exchange_calls"]), "exchange_ms_rank0": 1e3 * prof["exchange"],
        "phases_ms_rank0": {k: 1e3 * prof[k] for k in ("polycommit", "prove_sc_phase_one", "prove_sc_phase_two", "polyeval")},
        "proof_identical_on_all_ranks": all(d == digest for d in digests), "proof_identical_to_single_gpu": same_as_single,
    }


def main():
    claim_stdout()
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout to the one JSON line (NCCL prints its version banner there)
    if world == 1:
        os.environ.setdefault("SB200_BIND", "1")  # N = 1 binds its host threads next to the GPU exactly as the ranks of N > 1 do
    import torch
    import otti_b200 as sb
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    assert sb.lib().sb200_init(local_rank) == 0, sb.lib().sb200_last_error()
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    env = Env(args, torch, sb, dist, rank, world, local_rank)

    s = args.log_size
    n = 1 << s
    t0 = time.perf_counter()
    inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, NUM_INPUTS, SEED + rank, pinned=True)
    t_inst = time.perf_counter() - t0
    cold, cold_bytes = None, None
    if rank == 0:
        cold, cold_bytes, g1 = cold_path(env, n, inst, vars_, inputs, t_inst)
        del g1
    t0 = time.perf_counter()
    gens = sb.NIZKGens.new(n, n, NUM_INPUTS)
    torch.cuda.synchronize()
    gens_s = time.perf_counter() - t0

    proof = None
    for i in range(args.warmup):
        proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE)
    sampler = ClockSampler(local_rank)
    inst.upload_witness(gens, vars_)
    env.barrier()
    if rank == 0:
        sampler.start()
    res, e2e, launches, profs, profs_e2e, proof = timed_abab(env, inst, vars_, inputs, gens, args.steps)
    env.barrier()
    clocks = sampler.stop() if rank == 0 else None
    res_ms, e2e_ms = sum(res), sum(e2e)
    prof, prof_e2e = profs[-1], profs_e2e[-1]
    c = gens.window_bits

    t = torch.tensor([res_ms, e2e_ms], dtype=torch.float64, device="cuda")
    per_rank = None
    if dist is not None:
        mine = torch.tensor([res_ms / args.steps, e2e_ms / args.steps, float(sb.lib().sb200_bound_cpus())], dtype=torch.float64, device="cuda")
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = {"ms_per_step": [round(float(x[0]), 4) for x in allr], "e2e_ms_per_step": [round(float(x[1]), 4) for x in allr],
                    "host_cpus_bound": [int(x[2]) for x in allr]}
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res_ms, e2e_ms = float(t[0]), float(t[1])

    # which kernel dominates a step: per-kernel device time of two more proofs, every launch between CUDA events
    kp = kernel_profile(sb, lambda: sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE, resident=True), 2) if rank == 0 else []
    if cold is not None:
        cold["proof_same_as_steady_state"] = cold_bytes == proof.to_bytes()
    verify_ms = None
    if rank == 0:
        vt = []
        for _ in range(3):
            ta = time.perf_counter()
            proof.verify(inst, inputs, sb.Transcript.new(LABEL), gens)
            vt.append(1e3 * (time.perf_counter() - ta))
        verify_ms = min(vt)

    sharded = []
    if dist is not None:
        del inst, gens
        for tok in [x for x in args.shard_log_sizes.split(",") if x.strip()]:
            sharded.append(run_sharded(env, int(tok)))
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    step_ms = res_ms / args.steps
    value = world * n * args.steps / (res_ms * 1e-3)
    e2e_value = world * n * args.steps / (e2e_ms * 1e-3)
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0

    # ---- roofline of the kernel that dominates the step (chosen from the measured per-kernel times).  For the commitment
    # MSM the duration is the one measured live inside the timed steps (CUDA events around that launch on its stream).
    roofline = None
    if kp:
        dom, dcnt, dms = kp[0]
        if dom == "msm_fixed_kernel":
            k_s = sum(p["msm_commit_kernel"] for p in profs) / len(profs)
        else:
            k_s = dms * 1e-3 / max(dcnt, 1)
        roofline = roofline_for(dom, k_s, (k_s * (dcnt if dom != "msm_fixed_kernel" else 1)) / (step_ms * 1e-3), s, c, sm_mhz)
        roofline["chosen_by"] = "largest per-kernel device time of a profiled step (kernels_ms_per_proof below)"
        r0 = sum(p["sc_cubic_round0_kernel"] for p in profs) / len(profs)
        # round 0 of phase one reads Az, Bz, Cz once (eq(tau) is factored out of the round polynomial: kernels.cu sc_cubic_eq_kernel)
        roofline["hbm_kernel_sample"] = {"kernel": "sc_cubic_eq_kernel<false> (round 0: 3 tables x 32 B x 2^%d read; 4 F_l products per pair)" % s,
                                         "achieved": 3 * 32 * n / r0 / 1e9 if r0 > 0 else 0.0, "unit": "GB/s", "kernel_ms": 1e3 * r0,
                                         "int_pipe_frac": (4 * (n / 2) * 89 / r0) / int_pipe_peak(sm_mhz) if r0 > 0 else 0.0,
                                         "note": "bound by the integer pipe (89 IMAD.WIDE per F_l product), not by HBM"}

    cpu_baseline = None
    if not args.no_cpu_baseline and world == 1:
        import oracle_lib as O
        sb_s = min(s, 18)
        nb = 1 << sb_s
        oi, ov, oin = O.Instance.synthetic(nb, nb, NUM_INPUTS, SEED)
        og = O.Gens(nb, nb, NUM_INPUTS)
        t0 = time.perf_counter()
        want = O.nizk_prove(oi, ov, oin, og, LABEL, TAPE, 1)
        t1 = time.perf_counter()
        threads = host_threads()
        O.nizk_prove(oi, ov, oin, og, LABEL, TAPE, threads)
        t2 = time.perf_counter()
        parity = (want == proof.to_bytes()) if sb_s == s else None
        cpu_baseline = {
            "value": nb / (t1 - t0), "unit": "constraints/s", "cores": 1, "kind": "port",
            "sample": "1 full prove of synthetic R1CS 2^%d, CPU restatement of the reference (not curve25519-dalek, no AVX2 point backend)" % sb_s,
            "all_threads": {"value": nb / (t2 - t1), "cores": threads},
            "proof_bytes_equal_gpu": parity,
            "published_reference": {"value": README_2_20_CONSTRAINTS_PER_S, "unit": "constraints/s",
                                    "note": "Spartan/README.md:365-377, 2^20, 1 core i7-1065G7 (other hardware)"},
        }

    line = {
        "metric": METRIC, "value": value, "unit": "constraints/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None,  # BASELINE.md holds no published number for this metric on this hardware
        "dtype": DTYPE, "data": "synthetic",
        "config": {"workload": workload_name(s), "per_gpu": "one independent proof per GPU per step",
                   "l2": "256 MiB flush before every timed step", "order": "resident and end-to-end steps alternate (ABAB)",
                   "msm_window_bits": c, "host_threads": host_threads(), "gens_build_s_untimed": gens_s, "instance_build_s_untimed": t_inst,
                   "proof_bytes": len(proof.to_bytes())},
        "prove_time_s": step_ms * 1e-3,
        "step_ms": quantiles(res),
        "phases_ms": {k: 1e3 * sum(p[k] for p in profs) / len(profs)
                      for k in ("polycommit", "prove_sc_phase_one", "prove_sc_phase_two", "polyeval", "host_fiat_shamir")},
        "e2e": {"value": e2e_value, "unit": "constraints/s", "h2d_bytes_per_step": int(prof_e2e["h2d_bytes"]),
                "d2h_bytes_per_step": int(prof_e2e["d2h_bytes"]), "ms_per_step": e2e_ms / args.steps, "step_ms": quantiles(e2e)},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": roofline,
        "kernels_ms_per_proof": [{"kernel": k, "launches": cnt, "ms": round(ms, 4)} for k, cnt, ms in kp[:10]],
        "verify_ms": verify_ms,
        "cold": cold,
        "cpu_baseline": cpu_baseline,
        "published_reference": {"value": README_2_20_CONSTRAINTS_PER_S, "unit": "constraints/s", "verify_ms": 414.5,
                                "note": "Spartan/README.md:365-384, 2^20, 1 core i7-1065G7 (other hardware; not the same measurement)"},
    }
    if world == 1 and not args.no_side:
        del inst, gens
        side = {}
        try:
            side["synthetic_2p16"], snark = side_2p16(env)
            if snark is not None:
                line["snark_mode"] = snark
        except Exception as e:
            side["synthetic_2p16"] = {"error": "%s: %s" % (type(e).__name__, e)}
        if not args.no_snark:
            try:
                side["snark_2p20"] = side_snark_2p20(env)
            except Exception as e:
                side["snark_2p20"] = {"error": "%s: %s" % (type(e).__name__, e)}
        try:
            side["otti_shaped"] = side_otti(env, args.otti_log_size)
        except Exception as e:
            side["otti_shaped"] = {"error": "%s: %s" % (type(e).__name__, e)}
        try:
            side["otti_shaped_sweep"] = side_otti_sweep(env)
        except Exception as e:
            side["otti_shaped_sweep"] = {"error": "%s: %s" % (type(e).__name__, e)}
        line["side"] = side
    if per_rank is not None:
        line["per_rank"] = per_rank
    if sharded:
        line["sharded"] = sharded[0]
        if len(sharded) > 1:
            line["sharded_more"] = sharded[1:]
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
