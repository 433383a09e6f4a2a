#!/usr/bin/env python3
"""Rebuild the `data/` directory the reference's Catch2 tests read (tests/pfaai_tests.cpp:60-120) from the fixtures committed
under tests/golden/ -- the reference tree itself does not travel to the GPU box.

    python tests/binding/make_data.py OUTDIR        -> OUTDIR/data/...

  xdb_subset1.db, xdb_subset2.db, xdb_subset_combo12.db   gunzipped copies of the reference's own small databases
  modified_xantho_fastaai2.db    the reference's 20-genome database is absent from its repository (large blob); it is
                                 re-created here from its golden F / Lc arrays (xanthodb.npz: every posting (tetramer, SCP,
                                 genome) of the database) with the genome and SCP names of names.json.  The unmodified reference
                                 (oracle/_ref/pfaai_tests_ref.x) reproduces xanthodb_{jac,aji}.bin from it, which pins the copy.
  *_jac.bin, *_aji.bin, *_lc_array.bin, ...   the goldens, written back in cereal's BinaryOutputArchive layout
"""
import gzip
import json
import os
import shutil
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def lists_from_f(F, Lc, P, G):
    """Tetramer-major postings -> genome-major lists (seg_off over g*P+p, ascending tetramers)."""
    tet = np.repeat(np.arange(Lc.shape[0], dtype=np.int32), Lc)
    assert tet.shape[0] == F.shape[0]
    key = F["genome"].astype(np.int64) * P + F["protein"]
    order = np.argsort(key, kind="stable")                  # tetramers stay ascending inside a (genome, SCP) list
    off = np.zeros(G * P + 1, np.int64)
    np.cumsum(np.bincount(key, minlength=G * P), out=off[1:])
    return off, tet[order]


def main(out):
    from parfastaai_b200 import cereal_io, fastaai_db
    data = os.path.join(out, "data")
    os.makedirs(data, exist_ok=True)
    names = json.load(open(os.path.join(GOLDEN, "names.json")))
    for db in ("xdb_subset1", "xdb_subset2", "xdb_subset_combo12"):
        with gzip.open(os.path.join(GOLDEN, db + ".db.gz"), "rb") as src, open(os.path.join(data, db + ".db"), "wb") as dst:
            shutil.copyfileobj(src, dst)
    for f in ("qsub_test_input.txt", "qsub_test_bad_input.txt"):
        shutil.copy(os.path.join(GOLDEN, f), os.path.join(data, f))
    x = np.load(os.path.join(GOLDEN, "xanthodb.npz"))
    P, G = x["t"].shape
    off, vals = lists_from_f(x["f"], x["lc"], P, G)
    assert np.array_equal(np.diff(off).reshape(G, P).T, x["t"]), "T of the golden is not the list lengths of its F"
    fastaai_db.write_db(os.path.join(data, "modified_xantho_fastaai2.db"), names["xantho_genomes"], names["xantho_proteins"], off, vals)
    for pre, npz in (("xanthodb", "xanthodb"), ("xdb_subset1", "xdb_subset1"), ("xdb_subset2", "xdb_subset2"), ("xdb_qt", "xdb_qt"),
                     ("xdb_qry_subset", "xdb_qry_subset")):
        d = np.load(os.path.join(GOLDEN, npz + ".npz"))
        p = os.path.join(data, pre)
        if "lc" in d.files:
            cereal_io.write_int_vector(p + "_lc_array.bin", d["lc"])
            cereal_io.write_int_vector(p + "_lp_array.bin", d["lp"])
            cereal_io.write_f_array(p + "_f_array.bin", d["f"])
            cereal_io.write_t_matrix(p + "_t_matrix.bin", d["t"])
        if "sorted_e" in d.files:
            cereal_io.write_e_array(p + "_sorted_e_array.bin", d["sorted_e"])
        if "combo_t" in d.files:
            cereal_io.write_t_matrix(p + "_combo_t_matrix.bin", d["combo_t"])
        jac = np.zeros(d["jac"].shape[0], cereal_io.JAC_DTYPE)
        for f in ("gA", "gB", "S", "N"):
            jac[f] = d["jac"][f]
        cereal_io.write_jac(p + "_jac.bin", jac)
        cereal_io.write_double_vector(p + "_aji.bin", d["aji"])
    return data


if __name__ == "__main__":
    print(main(sys.argv[1] if len(sys.argv) > 1 else "."))
