#!/usr/bin/env python
"""bench.py -- Spartan NIZK prove throughput (constraints/s) on B200, per the driver contract.

A "step" is one NIZK::prove of one synthetic R1CS (Instance::produce_synthetic_r1cs,
r1csinstance.rs:136-214; 10 inputs, one non-zero per row and matrix, uniform 253-bit witness),
generators and instance built outside the timed region like benches/nizk.rs:27-39.

  value   witness already resident in HBM when the timed region starts
  e2e     the same prove through the public API (otti_b200.NIZK.prove) from PINNED host memory:
          host->device copy of the witness and device->host read of the proof inside the timed region
  N > 1   one process per GPU, each proving its own instance (independent proofs are the natural
          data-parallel unit of Otti's run.py loop over datasets); no data-path collective; weak scaling.

--impl reference times the CPU restatement of the reference prover (oracle/, "port": the real
curve25519-dalek prover is Rust and cannot be built in this image) with all host threads.
"""
import argparse
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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log-size", type=int, default=16, help="log2 of constraints = vars (configs[1] = 16)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-snark", action="store_true", help="skip the SNARK-mode side measurements (N = 1 only)")
    ap.add_argument("--shard-log-size", type=int, default=22,
                    help="N > 1: log2 size of the ONE proof additionally sharded over all N GPUs (configs[4]); 0 = skip")
    ap.add_argument("--exchange", default="shm", choices=["shm", "nccl"], help="exchange hook of the sharded proof")
    return ap.parse_args()


def snark_side_measurements(sb, inst, n, s):
    """The parts of SNARK mode that are built, on the bench's own instance, outside the timed region
    of the headline metric: whole-call milliseconds (best of 3 after one warm-up) of SNARK::encode, of the
    commit_nondet_witness step and of the build_layered_network step of SNARK::prove.  Never fails the bench line."""
    try:
        import random
        rng = random.Random(s)
        ell = 2**252 + 27742317777372353535851937790883648493
        sc = lambda k: b"".join(rng.randrange(ell).to_bytes(32, "little") for _ in range(k))
        t0 = time.perf_counter()
        gens = sb.SNARKGens.new(n, n, NUM_INPUTS, n)
        out = {"workload": "synthetic R1CS 2^%d (the bench instance), instance resident" % s,
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
        out["published_reference"] = {"encode_s": 14.26, "commit_nondet_witness_s": 14.45, "build_layered_network_s": 3.44,
                                      "note": "Spartan/README.md:305-345, 2^20, 1 core (other hardware)"}
        return out
    except Exception as e:  # a side measurement must not cost the headline line
        return {"error": "%s: %s" % (type(e).__name__, e)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md:70-72)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.p = gpu, None

    def start(self):
        try:
            def unbind():  # sb200_init bound this rank to its slice of the host cores: the sampler must not steal them
                try:
                    os.sched_setaffinity(0, range(os.cpu_count() or 1))
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
            return float(json.load(f)["hbm_gbs"]), "of measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "of fallback (B200_PROFILING.md)"


def run_reference(args, rank, world):
    """CPU restatement of the reference prover on the host cores (rank 0 only)."""
    if rank != 0:
        return
    import oracle_lib as O
    s = min(args.log_size, 18)  # bounded sample: throughput in constraints/s is size-independent to first order (O(N) prover)
    n = 1 << s
    threads = host_threads()
    inst, vars_, inputs = O.Instance.synthetic(n, n, NUM_INPUTS, SEED)
    gens = O.Gens(n, n, NUM_INPUTS)
    for _ in range(min(args.warmup, 1)):
        O.nizk_prove(inst, vars_, inputs, gens, LABEL, TAPE, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.nizk_prove(inst, vars_, inputs, gens, LABEL, TAPE, threads)
    dt = time.perf_counter() - t0
    value = n * args.steps / dt
    sample = "%d full proves of synthetic R1CS 2^%d on %d host threads (commitment rows threaded like the reference's `multicore` feature)" % (
        args.steps, s, threads)
    line = {
        "impl": "reference", "metric": "spartan_nizk_prove_constraints_per_s", "value": value, "unit": "constraints/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u256 (F_l 253-bit / F_p 255-bit integers)",
        "data": "synthetic", "config": {"workload": "synthetic Spartan R1CS 2^%d constraints = vars, 10 inputs, NIZK::prove" % args.log_size,
                                        "sample_log_size": s},
        "cpu_baseline": {"value": value, "unit": "constraints/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "constraints/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def run_sharded(args, sb, torch, dist, rank, world, local_rank, flush_l2, barrier):
    """ONE proof of 2^shard_log_size constraints sharded over all ranks: constraint rows / z columns
    cyclic, commitment rows in blocks, per-round exchange of 96 bytes per rank; every rank gets the same proof bytes.
    Time = max over ranks of the CUDA-event time around the prove.  Rank 0 then proves the same R1CS alone."""
    import hashlib
    import uuid
    from otti_b200 import comm as sbcomm
    s = args.shard_log_size
    n = 1 << s
    name = ["/sb200-bench-" + uuid.uuid4().hex[:12]]
    dist.broadcast_object_list(name, src=0)
    comm = sbcomm.ShmComm(name[0], rank, world) if args.exchange == "shm" else sbcomm.DistComm()
    inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, NUM_INPUTS, SEED, pinned=True, rank=rank, world=world)
    gens = sb.NIZKGens.new(n, n, NUM_INPUTS)
    proof = None
    for _ in range(args.warmup):
        proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE, comm=comm)
    ext = torch.cuda.ExternalStream(inst.stream)
    inst.upload_witness(gens, vars_)

    def timed(resident, c):
        nonlocal proof
        tot, prof = 0.0, None
        for i in range(args.steps):
            flush_l2(i)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ext)
            proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE, resident=resident, comm=c)
            e1.record(ext)
            e1.synchronize()
            tot += e0.elapsed_time(e1)
            prof = sb.NIZK.last_profile()
        return tot, prof

    res_ms, prof = timed(True, comm)
    e2e_ms, _ = timed(False, comm)
    digest = hashlib.sha256(proof.to_bytes()).digest()
    t = torch.tensor([res_ms, e2e_ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    digests = [None] * world
    dist.all_gather_object(digests, digest)
    comm.close()
    del inst
    single_ms = None
    if rank == 0:  # the same R1CS on one GPU, for the speed-up
        inst1, v1, i1 = sb.Instance.produce_synthetic_r1cs(n, n, NUM_INPUTS, SEED, pinned=True)
        for _ in range(args.warmup):
            p1 = sb.NIZK.prove(inst1, v1, i1, gens, sb.Transcript.new(LABEL), TAPE)
        ext1 = torch.cuda.ExternalStream(inst1.stream)
        inst1.upload_witness(gens, v1)
        tot = 0.0
        for i in range(args.steps):
            flush_l2(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ext1)
            p1 = sb.NIZK.prove(inst1, v1, i1, gens, sb.Transcript.new(LABEL), TAPE, resident=True)
            e1.record(ext1)
            e1.synchronize()
            tot += e0.elapsed_time(e1)
        single_ms = tot / args.steps
        same_as_single = hashlib.sha256(p1.to_bytes()).digest() == digest
    barrier()
    if rank != 0:
        return None
    ms = float(t[0]) / args.steps
    return {
        "workload": "ONE synthetic Spartan R1CS 2^%d sharded over %d GPUs (rows/columns cyclic, commitment rows in blocks)" % (s, world),
        "scaling": "strong", "prove_ms": ms, "constraints_per_s": n / (ms * 1e-3),
        "e2e_prove_ms": float(t[1]) / args.steps, "single_gpu_prove_ms": single_ms,
        "speedup_vs_single_gpu": (single_ms / ms) if single_ms else None,
        "exchange": args.exchange, "exchange_calls_per_proof": int(prof["exchange_calls"]), "exchange_ms_rank0": 1e3 * prof["exchange"],
        "phases_ms_rank0": {k: 1e3 * prof[k] for k in ("polycommit", "prove_sc_phase_one", "prove_sc_phase_two", "polyeval")},
        "proof_identical_on_all_ranks": all(d == digest for d in digests), "proof_identical_to_single_gpu": same_as_single,
    }


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


def main():
    claim_stdout()
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # keep stdout to the one JSON line (NCCL prints its version banner there)
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

    s = args.log_size
    n = 1 << s
    inst, vars_, inputs = sb.Instance.produce_synthetic_r1cs(n, n, NUM_INPUTS, SEED + rank, pinned=True)
    t0 = time.perf_counter()
    gens = sb.NIZKGens.new(n, n, NUM_INPUTS)
    torch.cuda.synchronize()
    gens_s = time.perf_counter() - t0

    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def flush_l2(i):
        flush_buf.fill_(i & 0xFF)
        torch.cuda.synchronize()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    proof = None
    for i in range(args.warmup):
        proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE)
    ext = torch.cuda.ExternalStream(inst.stream)

    def timed(resident):
        """K steps, each bracketed by CUDA events on the stream the kernels are launched on."""
        nonlocal proof
        total_ms, launches, prof = 0.0, 0, None
        for i in range(args.steps):
            flush_l2(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(ext)
            proof = sb.NIZK.prove(inst, vars_, inputs, gens, sb.Transcript.new(LABEL), TAPE, resident=resident)
            e1.record(ext)
            e1.synchronize()
            total_ms += e0.elapsed_time(e1)
            prof = sb.NIZK.last_profile()
            launches += int(prof["kernel_launches"])
        return total_ms, launches, prof

    sampler = ClockSampler(local_rank)
    inst.upload_witness(gens, vars_)
    barrier()
    if rank == 0:
        sampler.start()
    res_ms, launches, prof = timed(True)
    barrier()
    e2e_ms, _, prof_e2e = timed(False)
    barrier()
    clocks = sampler.stop() if rank == 0 else None

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

    sharded = None
    if dist is not None and args.shard_log_size > 0:
        sharded = run_sharded(args, sb, torch, dist, rank, world, local_rank, flush_l2, barrier)
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    value = world * n * args.steps / (res_ms * 1e-3)
    e2e_value = world * n * args.steps / (e2e_ms * 1e-3)

    # ---- roofline of the dominant kernel (commitment MSM), from CUDA events inside the timed region.
    # The kernel is bound by the integer-multiply pipe (fmaheavy), not by HBM (DESIGN.md "Rooflines"):
    #   algorithmic work  = terms x windows x (1 - 2^-c) mixed additions x 7 F_p multiplications x 75 IMAD.WIDE.U32
    #   algorithmic bytes = one 32-byte scalar per term + one 96-byte table entry per mixed addition
    ell = s
    Ls, Rs = 1 << (ell // 2), 1 << (ell - ell // 2)
    c = int(sb.lib().sb200_k_gens_window(gens.h))
    W = 252 // c + 1
    terms = n + Ls
    madds = terms * W * (1 - 2.0 ** -c)
    k_s = prof["msm_commit_kernel"]
    alg_bytes = terms * 32 + madds * 96
    hbm_peak, which = measured_peaks()
    achieved = alg_bytes / k_s / 1e9 if k_s > 0 else 0.0
    imad = madds * 7 * 75  # IMAD.WIDE.U32 per F_p product: 64 partial products + 8 (x38 fold) + 3 (cuobjdump, fp.cuh)
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    # fmaheavy issues one IMAD.WIDE.U32 warp instruction per 4 cycles per SM sub-partition (measured with
    # tools/microbench/fpmul_bench under ncu: 523 IMAD.WIDE per mixed addition, 2435 cycles at 85 % pipe-active) -> 8 lanes/clk/SMSP
    int_peak = 148 * 4 * 8 * sm_mhz * 1e6
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "r01_ncu_traffic.json")) as f:
            ent = json.load(f)["msm_fixed_kernel"].get(str(s))
        if ent:
            traffic, traffic_src = ent["bytes"], ent["source"]
    except Exception:
        pass
    roofline = {
        "kernel": "msm_fixed_kernel (commitment rows, %d x %d, %d-bit windows)" % (Ls, Rs, c), "bound": "hbm", "achieved": achieved,
        "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": traffic, "traffic_unit": "bytes per launch (ncu dram read + write)",
        "traffic_source": traffic_src, "algorithmic_bytes": alg_bytes, "peak_source": which,
        "kernel_ms": 1e3 * k_s, "share_of_step": k_s / (res_ms * 1e-3 / args.steps),
        "int_pipe": {"achieved_gimad_per_s": imad / k_s / 1e9 if k_s > 0 else 0.0, "peak_gimad_per_s": int_peak / 1e9,
                     "frac": (imad / k_s) / int_peak if k_s > 0 else 0.0,
                     "note": "the kernel's real limiter: 64-bit integer multiply-adds (IMAD.WIDE.U32) on the fmaheavy pipe; "
                             "ncu sm__pipe_fmaheavy_cycles_active 75.9 % at 2^22 (profiles/r01_ncu_full_2p22_v5.txt, r01_summary.md)"},
        "hbm_kernel_sample": {"kernel": "sc_cubic_kernel<false> (round 0: 4 tables x 32 B x 2^%d read)" % s,
                              "achieved": 4 * 32 * n / prof["sc_cubic_round0_kernel"] / 1e9 if prof["sc_cubic_round0_kernel"] > 0 else 0.0,
                              "unit": "GB/s", "kernel_ms": 1e3 * prof["sc_cubic_round0_kernel"]},
    }

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
        "metric": "spartan_nizk_prove_constraints_per_s", "value": value, "unit": "constraints/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": res_ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": (value / README_2_20_CONSTRAINTS_PER_S) if (s == 20 and world == 1) else None,
        "dtype": "u256 (F_l 253-bit / F_p 255-bit integers)", "data": "synthetic",
        "config": {"workload": "synthetic Spartan R1CS 2^%d constraints = vars, 10 inputs, NIZK::prove%s" % (
            s, " (configs[1])" if s == 16 else ""), "per_gpu": "one independent proof per GPU per step", "l2": "256 MiB flush between timed steps",
            "msm_window_bits": c, "host_threads": host_threads(), "gens_build_s_untimed": gens_s, "proof_bytes": len(proof.to_bytes())},
        "prove_time_s": res_ms * 1e-3 / args.steps,
        "phases_ms": {k: 1e3 * prof[k] for k in ("polycommit", "prove_sc_phase_one", "prove_sc_phase_two", "polyeval", "host_fiat_shamir")},
        "e2e": {"value": e2e_value, "unit": "constraints/s", "h2d_bytes_per_step": int(prof_e2e["h2d_bytes"]),
                "d2h_bytes_per_step": int(prof_e2e["d2h_bytes"]), "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": roofline,
        "cpu_baseline": cpu_baseline,
    }
    if world == 1 and not args.no_snark and s <= 20:
        line["snark_mode"] = snark_side_measurements(sb, inst, n, s)
    if per_rank is not None:
        line["per_rank"] = per_rank
    if sharded is not None:
        line["sharded"] = sharded
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
