#!/usr/bin/env python3
"""Fixture dumper: write Lc / Lp / F / T / JAC / AJI of a FastAAI database in the reference's cereal
BinaryOutputArchive layout, so that the reference's own test binary can be pointed at new goldens.

    python tools/dump_fixtures.py main.db PREFIX [-q query_list.txt | -r query.db] [--no-ref-compat]
    python tools/dump_fixtures.py --synthetic N [--scps P] [--seed S] out.db PREFIX     (also writes the database)

Writes PREFIX_lc_array.bin, PREFIX_lp_array.bin, PREFIX_f_array.bin, PREFIX_t_matrix.bin, PREFIX_jac.bin, PREFIX_aji.bin
(+ PREFIX_aji_matrix_wheader.csv through par_fastaai.x when it is built).  The arrays come from the host loader
(parfastaai_b200/fastaai_db.py) and the CUDA path through the C-ABI (a GPU is required; there is no CPU fallback)."""
import argparse
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("db")
    ap.add_argument("prefix")
    ap.add_argument("-q", "--query_subset")
    ap.add_argument("-r", "--query_db")
    ap.add_argument("--no-ref-compat", action="store_true")
    ap.add_argument("--synthetic", type=int, default=0, help="generate this many synthetic genomes into `db` first")
    ap.add_argument("--scps", type=int, default=122)
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--device", type=int, default=0)
    args = ap.parse_args()

    import parfastaai_b200 as pf
    from parfastaai_b200 import cereal_io, fastaai_db, synth

    if args.synthetic:
        off, vals = synth.generate(args.synthetic, args.scps, seed=args.seed)
        gn, sn = fastaai_db.synthetic_names(args.synthetic, args.scps)
        fastaai_db.write_db(args.db, gn, sn, off, vals, progress=True)
    if args.query_db:
        d = fastaai_db.load_query_target(args.db, args.query_db)
        mode, nq, qids, T = pf.QUERY_DB, len(d["qry_genomes"]), None, d["T"]
        flags = 0 if args.no_ref_compat or len(d["qry_genomes"]) != len(d["genomes"]) else pf.FLAG_REF_COMPAT_QT
        G = len(d["genomes"]) + nq
    else:
        d = fastaai_db.load_single(args.db)
        G, T, flags = len(d["genomes"]), None, 0
        if args.query_subset:
            names = open(args.query_subset).read().split()
            idx = {n: i for i, n in enumerate(d["genomes"])}
            missing = [n for n in names if n not in idx]
            if missing:
                sys.exit("query genomes missing from the database: " + ", ".join(missing))
            mode, qids, nq = pf.QUERY_SUBSET, np.array([idx[n] for n in names], np.int32), len(names)
        else:
            mode, qids, nq = pf.ALL_VS_ALL, None, 0
    P = len(d["proteins"])
    ctx = pf.PfaaiGpu(args.device)
    ctx.upload_index(P, G, d["F"], d["Lc"], T)
    ctx.build()
    arr = ctx.get_arrays(want_lists=False)
    res = ctx.compute(mode, qry_ids=qids, n_qry=nq, flags=flags, want_jac=True)
    ctx.close()
    pre = args.prefix
    cereal_io.write_int_vector(pre + "_lc_array.bin", arr["Lc"])
    cereal_io.write_int_vector(pre + "_lp_array.bin", arr["Lp"])
    cereal_io.write_f_array(pre + "_f_array.bin", d["F"])
    cereal_io.write_t_matrix(pre + "_t_matrix.bin", arr["T"])
    jac = np.zeros(res["jac"].shape[0], cereal_io.JAC_DTYPE)          # JACTuple is archived field by field: 20 B, no padding
    for f in ("gA", "gB", "S", "N"):
        jac[f] = res["jac"][f]
    cereal_io.write_jac(pre + "_jac.bin", jac)
    cereal_io.write_double_vector(pre + "_aji.bin", res["aji"])
    cli = os.path.join(ROOT, "parfastaai_b200", "par_fastaai.x")
    if os.access(cli, os.X_OK):
        cmd = [cli, args.db, pre + "_aji_matrix_wheader.csv", "--device", str(args.device)]
        if args.query_subset:
            cmd += ["-q", args.query_subset]
        if args.query_db:
            cmd += ["-r", args.query_db] + (["--no-ref-compat"] if args.no_ref_compat else [])
        subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL)
    print(f"wrote {pre}_{{lc_array,lp_array,f_array,t_matrix,jac,aji}}.bin: P={P} G={G} |F|={d['F'].shape[0]} pairs={res['aji'].shape[0]}")


if __name__ == "__main__":
    main()
