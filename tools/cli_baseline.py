#!/usr/bin/env python
"""BASELINE.md section 3, literally: the reference's unmodified command-line binary (oracle/_ref/rawalign -t <all cores>) on
multi-read FAST5 files, timed as a whole process, next to the compiled C++ consumer of our C ABI
(integration/rawalign_b200_map) on the same reads and the same index file; PAF lines are diffed (minus mt:f:).

  python tools/cli_baseline.py --config C2 --reads 4096 [--out profiles/r02_cli_C2.json]

TEST / MEASUREMENT INFRASTRUCTURE: uses oracle/_ref (the reference build) and the FAST5 writer in libra_refmap.so.
"""
import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

FLAGS = {"C1": ["-x", "viral", "--dtw-evaluate-chains", "--dtw-fill-method", "banded=0.1"],
         "C2": ["-x", "sensitive", "--dtw-evaluate-chains", "--dtw-border-constraint", "sparse", "--dtw-fill-method", "banded=0.1"],
         "C4": ["-x", "fast", "--dtw-evaluate-chains", "--dtw-border-constraint", "global", "--dtw-fill-method", "full"],
         "C5": ["-x", "fast", "--dtw-evaluate-chains", "--dtw-fill-method", "banded=0.15"]}
OURS = {"C1": ["--dtw-evaluate-chains", "--dtw-fill-method", "banded=0.1"],
        "C2": ["--dtw-evaluate-chains", "--dtw-border-constraint", "sparse", "--dtw-fill-method", "banded=0.1"],
        "C4": ["--dtw-evaluate-chains", "--dtw-border-constraint", "global", "--dtw-fill-method", "full"],
        "C5": ["--dtw-evaluate-chains", "--dtw-fill-method", "banded=0.15"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2", choices=list(FLAGS))
    ap.add_argument("--reads", type=int, default=4096)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    from oracle import RefMap
    from workload.mapbench import Dataset
    rm = RefMap()
    ds = Dataset(args.config, log=lambda *a: print(*a, file=sys.stderr))
    raw, off, ln, dac_off, dac_scale = ds.reads(args.reads, seed=2)
    from workload import synth
    d = tempfile.mkdtemp(prefix="ra_cli_")
    f5 = os.path.join(d, "fast5"); os.makedirs(f5)
    per = 4000                                                   # reads per multi-read FAST5
    for k, a in enumerate(range(0, args.reads, per)):
        b = min(args.reads, a + per)
        rm.write_fast5(os.path.join(f5, f"batch_{k:04d}.fast5"), raw, off[a:b], ln[a:b], first_id=a,
                       digitisation=synth.DAC_DIGITISATION, range_=synth.DAC_RANGE, offset=synth.DAC_OFFSET)
    n_thr = os.cpu_count() or 1
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "rawalign")
    paf_ref = os.path.join(d, "ref.paf")
    import resource
    cmd = [ref_bin] + FLAGS[args.config] + ["-t", str(n_thr), "-o", paf_ref, ds.index_path, f5]     # /usr/bin/time is not in this image: wall clock + getrusage
    t0 = time.perf_counter()
    r = subprocess.run(cmd, capture_output=True, text=True)
    t_ref = time.perf_counter() - t0
    if r.returncode != 0:
        raise SystemExit("reference CLI failed: " + r.stderr[-2000:])
    rss_kb = resource.getrusage(resource.RUSAGE_CHILDREN).ru_maxrss
    # ours: the compiled C++ consumer of the C ABI on the same samples (already extracted from the FAST5: the reader stays the reference's)
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "integration"), "rawalign_b200_map"], stdout=subprocess.DEVNULL)
    rb = os.path.join(d, "reads.bin")
    with open(rb, "wb") as f:
        f.write(np.uint32(args.reads).tobytes()); f.write(off.astype(np.uint64).tobytes()); f.write(ln.astype(np.uint32).tobytes())
        f.write(dac_off.astype(np.float32).tobytes()); f.write(dac_scale.astype(np.float32).tobytes())
        f.write(np.uint64(len(raw)).tobytes()); f.write(raw.astype(np.int16).tobytes())
    t0 = time.perf_counter()
    o = subprocess.run([os.path.join(ROOT, "integration", "rawalign_b200_map"), "--batch", "8192"] + OURS[args.config] + [ds.index_path, rb], capture_output=True, text=True)
    t_ours = time.perf_counter() - t0
    if o.returncode != 0:
        raise SystemExit("consumer failed: " + o.stderr[-2000:])
    strip = lambda l: re.sub(r"mt:f:[0-9.eE+-]+\t", "", l)
    ref_lines = {l.split("\t", 1)[0]: strip(l) for l in open(paf_ref).read().splitlines()}
    ours = o.stdout.splitlines()
    same = 0
    for k, l in enumerate(ours):
        nm = f"read_{k:08d}"                                     # read_id attribute our FAST5 writer gives read k
        want = ref_lines.get(nm)
        got = strip(l).replace(f"read_{k}\t", nm + "\t", 1)
        same += int(want == got)
    out = {"config": args.config, "reads": args.reads, "host_threads": n_thr,
           "reference_cli": {"cmd": " ".join(cmd), "wall_s": t_ref, "reads_per_s": args.reads / t_ref, "max_rss_kb": int(rss_kb),
                             "what": "whole program: FAST5 read + decode + index load + map loop + PAF printing"},
           "ours_consumer": {"wall_s": t_ours, "reads_per_s": args.reads / t_ours, "what": "whole program: CUDA context + index file load + reads.bin read + ra_map_batch + PAF printing"},
           "paf_lines_equal": same, "paf_lines": len(ours)}
    print(json.dumps(out))
    if args.out:
        with open(args.out, "w") as f:
            f.write(json.dumps(out, indent=1) + "\n")


if __name__ == "__main__":
    main()
