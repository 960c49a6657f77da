"""The CPU-baseline variants BASELINE.md section 3 names, on this box's host cores (one JSON line each):
the unmodified reference prover (libsnark::r1cs_gg_ppzksnark_prover through snark::generate_proof) on the
batch-8 aggregator statement, DEBUG off and DEBUG on (the reference's shipped default, CMakeLists.txt:98-102:
is_satisfied + a QAP re-evaluation inside every prove), on all cores and on 8 threads.

Usage: python scripts/cpu_variants.py [--batch 8] > profiles/r02_cpu_baseline_variants.jsonl"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from oracle import ref as R

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=8)
args = ap.parse_args()
cpu = bench.host_cpu()
for debug in (False, True):
    if not R.available(aggregator=True, debug=debug):
        print(json.dumps({"variant": "DEBUG on" if debug else "DEBUG off", "unavailable": "oracle/_ref build missing"}), flush=True)
        continue
    lib = R.Ref(aggregator=True, debug=debug)
    assert lib.session_aggregator(args.batch, 3) == 0
    t0 = time.perf_counter()
    lib.session_setup(9)
    t_setup = time.perf_counter() - t0
    sz = lib.session_sizes()
    for threads in sorted({os.cpu_count(), min(8, os.cpu_count())}, reverse=True):
        lib.set_threads(threads)
        ts = [lib.session_prove_stock() for _ in range(2)]
        assert min(ts) > 0, "reference proof did not verify"
        print(json.dumps({"variant": "DEBUG on" if debug else "DEBUG off", "threads": threads, "prove_s": min(ts),
                          "proofs_per_s": 1.0 / min(ts), "runs": ts, "setup_s": t_setup, "circuit": "aggregator batch %d" % args.batch,
                          "num_constraints": sz["num_constraints"], "m": sz["m"], "cpu": cpu,
                          "entry": "snark::generate_proof = libsnark::r1cs_gg_ppzksnark_prover (unmodified, oracle/_ref)"}), flush=True)
