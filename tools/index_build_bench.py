"""Row f4 measured: ra_index_build (ri_idx_gen on the device) against the reference's own indexer on the same FASTA.

    python tools/index_build_bench.py C2 [C4 ...]   ->  one JSON line per config on stdout

For each config of workload/mapbench.py: the reference's `rawalign -d` (all host threads; wall time includes FASTA parsing and
writing the index file), then ra_index_build from the same bases already in host memory (median of 3; includes the upload of the
bases and the allocation of the tables), and -- the parity check -- the exported table of the built index against the exported
table of the reference's index FILE loaded by ra_index_load_file, plus the reference mirror bit for bit."""
import json
import os
import subprocess
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import rawalign_b200 as ra
    from oracle import PRESETS
    from workload import synth
    from workload.mapbench import CONFIGS, Dataset, _ref_binary

    def dev_bytes(ptr, n):
        class _V:
            __cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        return torch.as_tensor(_V(), device="cuda:0")

    for name in sys.argv[1:] or ["C2"]:
        cfg = CONFIGS[name]
        ds = Dataset(name, log=lambda *a: print(*a, file=sys.stderr))
        e, q, lq, w, n = PRESETS[cfg["preset"]]
        out = {"config": name, "reference_bases": int(sum(cfg["ref"])), "n_sequences": len(cfg["ref"]), "preset": cfg["preset"]}
        rb = _ref_binary()
        if rb:
            synth.write_fasta(ds.fasta, ds.seqs); synth.write_pore_model(ds.model)
            tmp = ds.index_path + ".timing"
            t0 = time.perf_counter()
            r = subprocess.run([rb, "-x", cfg["preset"], "-t", str(os.cpu_count() or 4), "-p", ds.model, "-d", tmp, ds.fasta], capture_output=True, text=True)
            out["reference_indexer_s"] = time.perf_counter() - t0
            out["reference_indexer_threads"] = os.cpu_count()
            assert r.returncode == 0
            os.remove(tmp)
        pore = synth.pore_model_as_loaded()
        with ra.Context(0) as ctx:
            ts = []
            for _ in range(3):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                ix = ra.Index.build(ctx, ds.seqs, pore, e=e, q=q, lq=lq, k=6)
                torch.cuda.synchronize()
                ts.append(time.perf_counter() - t0)
                ix.close()
            out["ra_index_build_s"] = sorted(ts)[1]
            out["ra_index_build_runs_s"] = ts
            if os.environ.get("RA_INDEX_COMPARE", "1") == "1":
                with ra.Index.build(ctx, ds.seqs, pore, e=e, q=q, lq=lq, k=6) as ix:
                    got = ix.export()
                    ptr, nb = ix.tables()[2]
                    mirror = dev_bytes(ptr, nb).clone()
                    counts = ix.counts()
                t0 = time.perf_counter()
                with ra.Index.load_file(ctx, ds.index_path) as fx:
                    out["ra_index_load_file_s"] = time.perf_counter() - t0
                    want = fx.export()
                    ptr, nb = fx.tables()[2]
                    out["mirror_bit_identical"] = bool(torch.equal(mirror, dev_bytes(ptr, nb)))
                    out["table_identical"] = bool(all(np.array_equal(got[k], want[k]) for k in ("keys", "key_off", "vals")))
                    out["n_keys"], out["n_positions"] = int(counts[7]), int(counts[8])
                    out["index_file_by"] = ds.index_by
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
