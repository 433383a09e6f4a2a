"""FastAAI SQLite databases: host-side ingest (reference semantics) and a writer for synthetic DBs.

Host tooling around the GPU path -- "SQLite/tetramer loading stays on the host" (BASELINE north_star).
``load_single`` / ``load_query_target`` deliver exactly the arrays the reference's DataBaseInterface
delivers to its data-structure build (paths relative to the reference repo):

    protein order   SELECT DISTINCT scp_acc FROM scp_data          include/pfaai/db_helper.hpp:169-218
    genome order    SELECT genome_name FROM genome_metadata        include/pfaai/db_helper.hpp:59-107
    Lc              tetramerOccCounts                              include/pfaai/scp_db.hpp:121-158 (QT :402-448)
    F               proteinSetGPPairs, (tetramer, protein, genome) include/pfaai/scp_db.hpp:161-216 (QT :450-528)
    T               proteinTetramerCounts                          include/pfaai/scp_db.hpp:219-262 (QT :531-589)

One sequential pass per ``<SCP>_tetras`` table (no per-thread P-way UNION ALL); the
(tetramer, protein) ordering the reference gets from ``ORDER BY tetramer, source_table`` is produced by one
stable argsort.  ``write_db`` emits the schema of SURVEY App. A.1 so that the UNMODIFIED reference binary
(oracle/_ref) can be timed on synthetic inputs.
"""
from __future__ import annotations

import os
import sqlite3

import numpy as np

from ._lib import NTETRA, PG_DTYPE


def _names(con, schema=""):
    pre = f"{schema}." if schema else ""
    genomes = [r[0] for r in con.execute(f"SELECT genome_name FROM {pre}genome_metadata")]
    return genomes


def _proteins_single(con):
    return [r[0] for r in con.execute("SELECT DISTINCT scp_acc FROM scp_data")]


def _read_tetras(con, table, schema=""):
    """-> (tetramers int32 [rows], counts int32 [rows], genomes int32 [sum counts]) in tetramer order."""
    pre = f"{schema}." if schema else ""
    tets, cnts, blobs = [], [], []
    for t, blob in con.execute(f"SELECT tetramer, genomes FROM {pre}`{table}` ORDER BY tetramer"):
        tets.append(t)
        cnts.append(len(blob) // 4)
        blobs.append(blob)
    g = np.frombuffer(b"".join(blobs), dtype="<i4") if blobs else np.zeros(0, np.int32)
    return np.asarray(tets, np.int32), np.asarray(cnts, np.int32), g


def _assemble(per_protein):
    """per_protein: list over protein index of (tetramer-per-posting, genome-per-posting) arrays.
    -> Lc, F in the reference order (tetramer, protein, genome-as-stored)."""
    tet = np.concatenate([x[0] for x in per_protein]) if per_protein else np.zeros(0, np.int32)
    gen = np.concatenate([x[1] for x in per_protein]) if per_protein else np.zeros(0, np.int32)
    pro = np.concatenate([np.full(x[0].shape[0], p, np.int32) for p, x in enumerate(per_protein)]) \
        if per_protein else np.zeros(0, np.int32)
    Lc = np.bincount(tet, minlength=NTETRA).astype(np.int32)
    # postings were appended protein-major with tetramers ascending inside a protein: a stable sort on the
    # tetramer alone yields (tetramer, protein, blob order) == ORDER BY tetramer, source_table
    order = np.argsort(tet, kind="stable")
    F = np.empty(tet.shape[0], PG_DTYPE)
    F["protein"] = pro[order]
    F["genome"] = gen[order]
    return Lc, F


def load_single(path):
    """All-vs-all / query-subset input: dict(proteins, genomes, Lc, F, T)."""
    if not os.path.isfile(path):
        raise FileNotFoundError(path)
    con = sqlite3.connect(f"file:{path}?mode=ro", uri=True)
    try:
        proteins = _proteins_single(con)
        genomes = _names(con)
        P, G = len(proteins), len(genomes)
        per = []
        T = np.zeros((P, G), np.int32)
        for p, acc in enumerate(proteins):
            tets, cnts, gen = _read_tetras(con, f"{acc}_tetras")
            per.append((np.repeat(tets, cnts), gen))
            for gid, nbytes in con.execute(f"SELECT genome_id, length(tetramers) FROM `{acc}_genomes`"):
                T[p, gid] = nbytes // 4
        Lc, F = _assemble(per)
    finally:
        con.close()
    return {"proteins": proteins, "genomes": genomes, "Lc": Lc, "F": F, "T": T}


def load_query_target(main_path, query_path):
    """Query-DB mode (-r): targets = main DB genomes (ids 0..nT-1), queries = query DB genomes (ids nT+q).
    SCPs = those in BOTH DBs; Lc/F restricted to tetramers present in both DBs for the SCP (inner join),
    T unrestricted.  dict(proteins, genomes (targets), qry_genomes, Lc, F, T [P x (nT+nQ)])."""
    for p in (main_path, query_path):
        if not os.path.isfile(p):
            raise FileNotFoundError(p)
    con = sqlite3.connect(f"file:{main_path}?mode=ro", uri=True)
    try:
        con.execute("ATTACH DATABASE ? AS QueryDB", (f"file:{query_path}?mode=ro",))
        proteins = [r[0] for r in con.execute(
            "SELECT DISTINCT a.scp_acc FROM scp_data a INNER JOIN QueryDB.scp_data b ON a.scp_acc = b.scp_acc")]
        genomes = _names(con)
        qry = _names(con, "QueryDB")
        nT, nQ = len(genomes), len(qry)
        P = len(proteins)
        T = np.zeros((P, nT + nQ), np.int32)
        per = []
        for p, acc in enumerate(proteins):
            tt, tc, tg = _read_tetras(con, f"{acc}_tetras")
            qt, qc, qg = _read_tetras(con, f"{acc}_tetras", "QueryDB")
            common, it, iq = np.intersect1d(tt, qt, assume_unique=True, return_indices=True)
            toff = np.concatenate([[0], np.cumsum(tc)])
            qoff = np.concatenate([[0], np.cumsum(qc)])
            tets, gens = [], []
            for k in range(common.shape[0]):   # per tetramer: target genomes first, then query ids + nT
                a, b = it[k], iq[k]
                tets.append(np.full(tc[a] + qc[b], common[k], np.int32))
                gens.append(tg[toff[a]:toff[a + 1]])
                gens.append(qg[qoff[b]:qoff[b + 1]] + nT)
            per.append((np.concatenate(tets) if tets else np.zeros(0, np.int32),
                        np.concatenate(gens).astype(np.int32) if gens else np.zeros(0, np.int32)))
            for gid, nbytes in con.execute(f"SELECT genome_id, length(tetramers) FROM `{acc}_genomes`"):
                T[p, gid] += nbytes // 4
            for gid, nbytes in con.execute(f"SELECT genome_id, length(tetramers) FROM QueryDB.`{acc}_genomes`"):
                T[p, nT + gid] += nbytes // 4
        Lc, F = _assemble(per)
    finally:
        con.close()
    return {"proteins": proteins, "genomes": genomes, "qry_genomes": qry, "Lc": Lc, "F": F, "T": T}


def write_db(path, genome_names, scp_names, seg_off, vals, progress=False):
    """Write genome-major tetramer lists (list (g,p) = vals[seg_off[g*P+p] : seg_off[g*P+p+1]], ascending)
    as a FastAAI database readable by the unmodified reference."""
    G, P = len(genome_names), len(scp_names)
    seg_off = np.asarray(seg_off, np.int64)
    vals = np.asarray(vals, np.int32)
    if os.path.exists(path):
        os.remove(path)
    con = sqlite3.connect(path)
    con.execute("PRAGMA journal_mode=OFF")
    con.execute("PRAGMA synchronous=OFF")
    cur = con.cursor()
    cur.execute("CREATE TABLE genome_metadata (genome_name TEXT, genome_id INTEGER PRIMARY KEY, genome_length INTEGER, "
                "genome_class INTEGER, SCP_count INTEGER)")
    cur.execute("CREATE TABLE scp_data (genome_id INTEGER, SCP_acc TEXT, SCP_score REAL, tetra_count INTEGER)")
    cur.executemany("INSERT INTO genome_metadata VALUES (?,?,?,?,?)",
                    [(n, g, 0, 0, P) for g, n in enumerate(genome_names)])
    lens = np.diff(seg_off).reshape(G, P)
    for p, acc in enumerate(scp_names):     # SCP-major so that SELECT DISTINCT returns scp_names order
        cur.executemany("INSERT INTO scp_data VALUES (?,?,?,?)", ((g, acc, 0.0, int(lens[g, p])) for g in range(G)))
    # one SCP at a time (streaming: memory beyond the input lists is O(postings of one SCP), so 50,000-genome databases --
    # SURVEY 8(f) rank 4 -- are written in one pass), everything per SCP vectorised: gather the SCP's lists, one stable argsort
    gidx = np.arange(G, dtype=np.int64)
    for p, acc in enumerate(scp_names):
        cur.execute(f"CREATE TABLE `{acc}_tetras` (tetramer INTEGER PRIMARY KEY, genomes BLOB)")
        cur.execute(f"CREATE TABLE `{acc}_genomes` (genome_id INTEGER PRIMARY KEY, tetramers BLOB)")
        beg, ln = seg_off[gidx * P + p], lens[:, p]
        ends = np.cumsum(ln)
        starts = ends - ln
        # position k of the concatenated lists of this SCP -> index into vals
        src = np.repeat(beg - starts, ln) + np.arange(int(ends[-1]) if G else 0, dtype=np.int64)
        tets = vals[src].astype("<i4")
        gens = np.repeat(np.arange(G, dtype=np.int32), ln)
        raw = tets.tobytes()
        cur.executemany(f"INSERT INTO `{acc}_genomes` VALUES (?,?)",
                        ((g, raw[4 * int(starts[g]):4 * int(ends[g])]) for g in range(G)))
        order = np.argsort(tets, kind="stable")           # genomes stay ascending inside a tetramer
        tets, gens = tets[order], gens[order].astype("<i4")
        uniq, start = np.unique(tets, return_index=True)
        end = np.append(start[1:], tets.shape[0])
        graw = gens.tobytes()
        cur.executemany(f"INSERT INTO `{acc}_tetras` VALUES (?,?)",
                        ((int(uniq[i]), graw[4 * int(start[i]):4 * int(end[i])]) for i in range(uniq.shape[0])))
        cur.execute(f"CREATE INDEX `{acc}_tetras_index` ON `{acc}_tetras` (tetramer)")
        if progress and (p + 1) % 10 == 0:
            print(f"[write_db] {p + 1}/{P} SCPs", flush=True)
    con.commit()
    con.close()
    return path


def synthetic_names(N, P):
    return [f"syn_genome_{g:06d}" for g in range(N)], [f"SYN{p:05d}.1" for p in range(P)]
