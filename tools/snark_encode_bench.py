"""SNARK::encode (the computation commitment, lib.rs:321-333) on the B200 against the CPU restatement.

    python tools/snark_encode_bench.py [--log-sizes 16 20] [--cpu-max-log 16]

Prints one JSON line per size: device milliseconds of the dense representation and of the two commitments (CUDA events
inside the library), the whole call on the host clock (allocation and the device->host copy of the commitments
included; the instance is resident, as it is for the reference's SNARK::encode(&inst, &gens)), and -- up to
--cpu-max-log -- the oracle's time for the same call with one thread and with every host core.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log-sizes", type=int, nargs="+", default=[16, 20])
    ap.add_argument("--cpu-max-log", type=int, default=16)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    import otti_b200 as sb
    assert sb.lib().sb200_init(0) == 0
    for s in args.log_sizes:
        n = 1 << s
        inst = sb.Instance.produce_synthetic_r1cs(n, n, 10, seed=0x0771)[0]
        t0 = time.perf_counter()
        gens = sb.SNARKGens.new(n, n, 10, n)
        t_gens = time.perf_counter() - t0
        best = None
        for _ in range(args.reps + 1):  # first call warms up
            cc = sb.SNARK.encode(inst, gens)
            p = sb.SNARK.last_encode_profile()
            if best is None or p["total_ms"] < best["total_ms"]:
                best = p
        line = {"op": "SNARK::encode", "workload": "synthetic R1CS 2^%d (3 x 2^%d non-zero entries)" % (s, s),
                "terms_comb_ops": 16 * n, "terms_comb_mem": 4 * n, "gens_build_s_untimed": round(t_gens, 3),
                "commitment_bytes": len(cc.bytes), **{k: round(v, 4) for k, v in best.items()}}
        if s <= args.cpu_max_log:
            import oracle_lib as O
            oinst = O.Instance.synthetic(n, n, 10, 0x0771)[0]
            for th, key in ((1, "cpu_1_thread_ms"), (os.cpu_count() or 1, "cpu_all_threads_ms")):
                t0 = time.perf_counter()
                ob = O.snark_encode(oinst, n, threads=th)
                line[key] = round((time.perf_counter() - t0) * 1e3, 1)
                assert ob == cc.bytes, "CUDA path and oracle disagree"
            line["cpu_threads"] = os.cpu_count() or 1
            line["bytes_equal_oracle"] = True
        print(json.dumps(line), flush=True)
        # commit_nondet_witness of SparseMatPolyEvalProof::prove at a random point
        import random
        rng = random.Random(s)
        ELL = 2**252 + 27742317777372353535851937790883648493
        rx = b"".join(rng.randrange(ELL).to_bytes(32, "little") for _ in range(s))
        ry = b"".join(rng.randrange(ELL).to_bytes(32, "little") for _ in range(s + 1))
        best = None
        for _ in range(args.reps + 1):
            pts = sb.SNARK.commit_derefs(inst, gens, rx, ry)
            p = sb.SNARK.last_encode_profile()
            if best is None or p["total_ms"] < best["total_ms"]:
                best = p
        line = {"op": "commit_nondet_witness (Derefs::commit)", "workload": "synthetic R1CS 2^%d" % s, "terms": 8 * n,
                "eq_dense_gather_ms": round(best["dense_ms"], 4), "commit_ms": round(best["commit_ops_ms"], 4),
                "total_ms": round(best["total_ms"], 4)}
        if s <= args.cpu_max_log:
            for th, key in ((1, "cpu_1_thread_ms"), (os.cpu_count() or 1, "cpu_all_threads_ms")):
                t0 = time.perf_counter()
                ob = O.snark_commit_derefs(oinst, n, rx, ry, threads=th)
                line[key] = round((time.perf_counter() - t0) * 1e3, 1)
                assert ob == b"".join(pts), "CUDA path and oracle disagree"
            line["bytes_equal_oracle"] = True
        print(json.dumps(line), flush=True)
        # build_layered_network of SparseMatPolyEvalProof::prove
        rh = rng.randrange(ELL).to_bytes(32, "little")
        rm = rng.randrange(ELL).to_bytes(32, "little")
        best = None
        for _ in range(args.reps + 1):
            prods = sb.SNARK.network_products(inst, rx, ry, rh, rm)
            p = sb.SNARK.last_encode_profile()
            if best is None or p["total_ms"] < best["total_ms"]:
                best = p
        line = {"op": "build_layered_network (PolyEvalNetwork::new)", "workload": "synthetic R1CS 2^%d" % s,
                "hashed_entries": 2 * (4 * n + 6 * n), "eq_dense_ms": round(best["dense_ms"], 4),
                "hash_layers_ms": round(best["commit_ops_ms"], 4), "product_circuits_ms": round(best["commit_mem_ms"], 4),
                "total_ms": round(best["total_ms"], 4)}
        if s <= args.cpu_max_log:
            t0 = time.perf_counter()
            ob = O.snark_network_products(oinst, rx, ry, rh, rm)
            line["cpu_1_thread_ms"] = round((time.perf_counter() - t0) * 1e3, 1)
            assert ob == prods, "CUDA path and oracle disagree"
            line["bytes_equal_oracle"] = True
        print(json.dumps(line), flush=True)
        del gens, inst


if __name__ == "__main__":
    main()
