#!/usr/bin/env python3
"""Regenerates tests/golden/ from the reference's own fixtures.

Run in the build container (where /root/reference is mounted):

    python tests/golden/make_golden.py [/root/reference]

The reference's golden vectors (``data/*.bin``, cereal binary archives; layouts in
``parfastaai_b200/cereal_io.py``) are re-packed, value for value, into compressed
``.npz`` files so that the GPU box (which has no /root/reference) can run the parity
tests.  CSV goldens, the query list and the three shipped SQLite DBs are stored
verbatim (DBs gzip'ed).  Name tables come from ``tests/pfaai_tests.hpp`` macros.

Nothing here is reference *source*: only data fixtures and name lists.
"""
import gzip
import json
import os
import re
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from parfastaai_b200 import cereal_io as cio  # noqa: E402


def macro_strings(text, name):
    m = re.search(r"#define\s+" + name + r"\s+\\?\s*\{(.*?)\}", text, re.S)
    if not m:
        raise KeyError(name)
    return re.findall(r'"([^"]+)"', m.group(1))


def main():
    ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
    data = os.path.join(ref, "data")
    out = HERE

    def D(name):
        return os.path.join(data, name)

    def pack(prefix, has_e=True, extra=None):
        d = {
            "lc": cio.read_int_vector(D(f"{prefix}_lc_array.bin")),
            "lp": cio.read_int_vector(D(f"{prefix}_lp_array.bin")),
            "f": cio.read_f_array(D(f"{prefix}_f_array.bin")),
            "t": cio.read_t_matrix(D(f"{prefix}_t_matrix.bin")),
            "jac": cio.read_jac(D(f"{prefix}_jac.bin")),
            "aji": cio.read_double_vector(D(f"{prefix}_aji.bin")),
        }
        if has_e:
            d["sorted_e"] = cio.read_e_array(D(f"{prefix}_sorted_e_array.bin"))
        if extra:
            d.update(extra)
        np.savez_compressed(os.path.join(out, f"{prefix}.npz"), **d)
        print(prefix, {k: v.shape for k, v in d.items()})

    # C1: 20-genome xantho DB (sorted-E fixture is missing upstream: .MISSING_LARGE_BLOBS)
    pack("xanthodb", has_e=False, extra={
        "gpe_starts": cio.read_int_vector(D("xanthodb_gpe_starts.bin")),
        "gpe_ends": cio.read_int_vector(D("xanthodb_gpe_ends.bin")),
        "e_size": cio.read_int_vector(D("xanthodb_e_size.bin")),
        "e_starts": cio.read_int_vector(D("xanthodb_e_starts.bin")),
    })
    pack("xdb_subset1")
    pack("xdb_subset2")
    pack("xdb_qt", extra={"combo_t": cio.read_t_matrix(D("xdb_qt_combo_t_matrix.bin"))})
    # C2: query-subset JAC/AJI (5 queries x 20 genomes, 85 pairs), same Lc/Lp/F/T as C1
    np.savez_compressed(os.path.join(out, "xdb_qry_subset.npz"),
                        jac=cio.read_jac(D("xdb_qry_subset_jac.bin")),
                        aji=cio.read_double_vector(D("xdb_qry_subset_aji.bin")))

    for name in ["xanthodb_aji_matrix_wheader.csv", "xanthodb_aji_matrix.csv",
                 "xdb_subset1_aji_matrix_wheader.csv", "xdb_subset2_aji_matrix_wheader.csv",
                 "qsub_test_output_matrix_wheader.csv", "qsub_test_input.txt",
                 "qsub_test_bad_input.txt"]:
        shutil.copyfile(D(name), os.path.join(out, name))
    for name in ["xdb_subset1.db", "xdb_subset2.db", "xdb_subset_combo12.db"]:
        with open(D(name), "rb") as fi, gzip.GzipFile(os.path.join(out, name + ".gz"), "wb",
                                                       compresslevel=9, mtime=0) as fo:
            shutil.copyfileobj(fi, fo)

    hdr = open(os.path.join(ref, "tests", "pfaai_tests.hpp")).read()
    names = {
        "xantho_proteins": macro_strings(hdr, "TESTDB_PROTEIN_SET"),
        "qt_proteins": macro_strings(hdr, "QT_TESTDB_PROTEIN_SET"),
        "xantho_genomes": macro_strings(hdr, "TESTDB_GENOME_SET"),
        "subset1_genomes": macro_strings(hdr, "TESTDB_SUBSET1_GENOME_SET"),
        "subset2_genomes": macro_strings(hdr, "TESTDB_SUBSET2_GENOME_SET"),
    }
    with open(os.path.join(out, "names.json"), "w") as fh:
        json.dump(names, fh, indent=1)
    print({k: len(v) for k, v in names.items()})


if __name__ == "__main__":
    main()
