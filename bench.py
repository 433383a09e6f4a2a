#!/usr/bin/env python3
"""bench.py -- genome-pair AJI/sec of the B200 AJI path (BASELINE.json metric), one JSON line on rank 0.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--genomes G] [--scps P]

A *step* is one pass of the hot path over one synthetic database (SURVEY 8(d) recipe, seed 1234):
device index build (Lp scan, (genome,SCP) histogram = T, stable radix sort of the postings into
genome-major lists) + row-slice packing + the tile kernel (counts -> Jaccard -> S,N -> AJI matrix).

  value   pairs/s with the reference-format index (F, Lc) already resident in HBM when the timed region
          starts and the AJI matrix left in HBM (CUDA events on the launching stream, max over ranks).
  e2e     the same metric through the host-buffer C-ABI calls a reference-side binding makes
          (pfaai_gpu_upload_index + pfaai_gpu_build + pfaai_gpu_compute):
          pinned host F/Lc copied H2D and the fp64 matrix copied D2H inside the timed region, every step.
  roofline     the tile kernel: algorithmic bytes / its CUDA-event time, against the
               measured HBM copy bandwidth in MEASURED_PEAKS.json.
  cpu_baseline the UNMODIFIED reference binary (oracle/_ref, kind "reference") on a bounded synthetic
               SQLite DB with all host cores; falls back to the oracle port when the binary is unusable.

Workload at every N: BASELINE config 5 (50,000 genomes x 122 SCPs all-vs-all, the configuration the metric is quoted on at
1/2/4/8 B200; it fits one GPU), output row blocks sharded over the ranks: strong scaling, the driver's per-N values are
directly comparable.  `--genomes 10000` gives BASELINE config 4.
`--impl reference` times the reference's own CPU implementation (rank 0 only).
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "genome-pair AJI/sec"
UNIT = "pairs/s"
SEED = 1234
REF_PHASES = ["Lc & Lp contruction", "F contruction", "T construction", "E constr.   (fin)", "JAC Construction"]


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled DURING the timed region
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def __enter__(self):
        try:
            fd, self.path = tempfile.mkstemp(prefix="pfaai_clk_", suffix=".csv")
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=fd, stderr=subprocess.DEVNULL)
            os.close(fd)
            time.sleep(0.3)
        except Exception:
            self.proc = None
        return self

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if not self.path or not os.path.exists(self.path):
            return out
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        with open(self.path) as fh:
            for line in fh:
                f = [x.strip() for x in line.split(",")]
                if len(f) < 7:
                    continue
                try:
                    sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
                except ValueError:
                    continue
                for n, v in zip(names, f[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
        os.remove(self.path)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), power_w_max=float(max(pw)),
                       reasons=sorted(reasons), samples=len(sm))
        return out


# ---------------------------------------------------------------------------------------------------
# reference arm: the unmodified reference binary on a synthetic FastAAI SQLite DB
# ---------------------------------------------------------------------------------------------------
def ref_binary():
    p = os.path.join(ROOT, "oracle", "_ref", "par_fastaai_ref.x")
    return p if os.path.exists(p) and os.access(p, os.X_OK) else None


def make_ref_db(n_genomes, P, tmpdir):
    from parfastaai_b200 import fastaai_db, synth
    off, vals = synth.generate(n_genomes, P, seed=SEED)
    gn, sn = fastaai_db.synthetic_names(n_genomes, P)
    path = os.path.join(tmpdir, f"syn_{n_genomes}x{P}.db")
    fastaai_db.write_db(path, gn, sn, off, vals)
    return path


def run_ref_once(db, threads, tmpdir):
    """-> (pairs/s over the reference's own phase timers, dict of phase ms, wall seconds)."""
    out_csv = os.path.join(tmpdir, "ref_out.csv")
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    t0 = time.perf_counter()
    r = subprocess.run([ref_binary(), db, out_csv], capture_output=True, text=True, env=env)
    wall = time.perf_counter() - t0
    if r.returncode != 0:
        raise RuntimeError(f"reference binary rc={r.returncode}: {r.stderr[-300:]}")
    phases = {}
    for line in r.stdout.splitlines():
        m = re.match(r"^\s*(.+?)\s*:\s*([0-9.eE+\-]+)\s*ms", line)
        if m:
            phases[m.group(1).strip()] = float(m.group(2))
    missing = [p for p in REF_PHASES if p.strip() not in phases]
    if missing:
        raise RuntimeError(f"reference stdout lacks phases {missing}")
    return phases, wall


REF_PATH_PHASES = ["E constr.   (fin)", "JAC Construction"]


def ref_path_seconds(phases):
    # the part of construct() + run() the GPU arm covers: from Lc / F / T in host memory to JAC / AJI, i.e. E generation + sort and
    # computeJAC / computeAJI (interface.hpp:319-323, algorithm_impl.hpp:281-329).  The SQLite-facing phases (Lc & Lp, F, T
    # construction) are parsed and reported, not counted: the GPU arm starts from host arrays too (like for like).
    return sum(phases[p.strip()] for p in REF_PATH_PHASES) / 1e3


def ref_ingest_seconds(phases):
    return sum(phases[p.strip()] for p in REF_PHASES if p not in REF_PATH_PHASES) / 1e3


def port_baseline(P, off, vals, T, n_rows, n_cols, gpu_matrix=None):
    """The oracle's OpenMP definition-level port on a rectangle of the bench workload (kind "port")."""
    from oracle import pfaai_oracle as orc
    G = T.shape[1]
    rows = np.arange(min(n_rows, G), dtype=np.int32)
    cols = np.arange(G - min(n_cols, G), G, dtype=np.int32)
    t0 = time.perf_counter()
    d = orc.direct_aji(P, G, off, vals, T, rows, cols)
    dt = time.perf_counter() - t0
    pairs = int(rows.shape[0]) * int(cols.shape[0]) - int(np.intersect1d(rows, cols).shape[0])
    res = {"value": pairs / dt, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
           "sample": f"oracle pfo_direct_aji (OpenMP sorted-list merge) on rows 0..{rows.shape[0]-1} x cols {cols[0]}..{cols[-1]} "
                     f"of the bench workload, {pairs} pairs in {dt:.2f} s"}
    if gpu_matrix is not None:
        sub = gpu_matrix[np.ix_(rows, cols)]
        res["max_abs_diff_vs_gpu"] = float(np.abs(sub - d["aji"]).max())
    return res


def verify_rectangles(G, rank, world):
    """Rectangles (name, rows, cols) of THIS rank's share of the all-vs-all matrix that the oracle re-computes after the timed
    steps: a diagonal tile (triangular skipping, zero diagonal, both triangles), a tile whose column dictionary was built by
    another rank (it travelled through the exchange), the ragged last row block and the last column block."""
    from parfastaai_b200 import sharding
    B = sharding.BLOCK
    nb = (G + B - 1) // B
    mine = [b for b in range(nb) if sharding.owner(b, world) == rank]
    rects = []
    if not mine:
        return rects
    I0 = mine[0]
    r0 = I0 * B
    rects.append(("diagonal", np.arange(r0, min(r0 + 24, G)), np.arange(r0, min(r0 + 96, G))))
    # a row block of ours with a later row block (its column dictionaries) owned by someone else
    late = None
    for I in reversed(mine):
        cands = [J for J in range(I + 1, nb) if world == 1 or sharding.owner(J, world) != rank]
        if cands:
            late = (I, cands[len(cands) // 2])
            break
    if late is not None:
        I, J = late
        ra = min(I * B + 200, G - 1)
        ca = min(J * B + 40, G - 1)
        rects.append(("late_list", np.arange(ra, min(ra + 24, (I + 1) * B, G)), np.arange(ca, min(ca + 112, G))))
    if sharding.owner(nb - 1, world) == rank:
        rects.append(("ragged_last_row_block", np.arange(max(0, G - 16), G), np.arange(max(0, G - 150), G)))
    ra = min(r0 + 100, G - 1)
    rects.append(("last_column_block", np.arange(ra, min(ra + 16, G)), np.arange(max(0, G - 64), G)))
    return [(n, r.astype(np.int32), c.astype(np.int32)) for n, r, c in rects if r.size and c.size]


def verify_against_oracle(G, P, dense, rects, read_cells, lower_valid):
    """Re-generate the lists of the rectangles' genomes (every synthetic list is a pure function of seed, G, genome, SCP), run
    the oracle's definition-level port on them and compare with the cells the GPU path stored.  read_cells(rows, cols) returns
    the stored values; lower_valid = cells with row genome > column genome were stored too (mirror images)."""
    from oracle import pfaai_oracle as orc
    from parfastaai_b200 import synth
    out = {"max_abs_diff": 0.0, "cells": 0, "rects": []}
    if not rects:
        return out
    ids = np.unique(np.concatenate([np.concatenate([r, c]) for _n, r, c in rects])).astype(np.int32)
    pos = {int(g): k for k, g in enumerate(ids)}
    off, vals = synth.generate_subset(G, P, SEED, ids, dense=dense)
    T = synth.t_matrix(ids.shape[0], P, off)
    for name, rows, cols in rects:
        lr = np.array([pos[int(g)] for g in rows], np.int32)
        lc = np.array([pos[int(g)] for g in cols], np.int32)
        want = orc.direct_aji(P, ids.shape[0], off, vals, T, lr, lc)["aji"]
        got = read_cells(rows, cols)
        mask = np.ones(want.shape, bool) if lower_valid else (rows[:, None] < cols[None, :])
        d = float(np.abs(got - want)[mask].max()) if mask.any() else 0.0
        if lower_valid and name == "late_list":      # the mirror images M[j][i] of an off-diagonal tile
            d = max(d, float(np.abs(read_cells(cols, rows) - want.T).max()))
        if "hit_rate" not in out and cols.shape[0] >= 64 and not np.intersect1d(rows, cols).size:
            # by-product for the on-chip ceiling: the fraction of a row's tetramers found in a 64-column dictionary
            hits = tot = 0
            for p in range(P):
                blk = np.unique(np.concatenate([vals[off[k * P + p]:off[k * P + p + 1]] for k in lc[:64]]))
                for k in lr:
                    x = vals[off[k * P + p]:off[k * P + p + 1]]
                    hits += int(np.isin(x, blk, assume_unique=True).sum()); tot += int(x.shape[0])
            out["hit_rate"] = hits / max(tot, 1)
        out["max_abs_diff"] = max(out["max_abs_diff"], d)
        out["cells"] += int(mask.sum())
        out["rects"].append({"name": name, "rows": [int(rows[0]), int(rows[-1])], "cols": [int(cols[0]), int(cols[-1])],
                             "max_abs_diff": d})
    return out


def dense_family_record(pf, synth, device, P, genomes=10000, reps=3):
    """Untimed side record of the N = 1 line: the real-like (dense) synthetic family at BASELINE config-4 size, all-vs-all, on the
    list path and on the path the cost model picks (the tensor-core contraction, aji_mma_kernel).  Tile-kernel time of both
    (CUDA events inside the C-ABI, best of `reps` after one warm-up), the plan, and an oracle check of the picked path's cells.
    Never part of `value` / `e2e`; it exists so that the driver's own run exercises and checks the dense path."""
    t0 = time.perf_counter()
    off, vals = synth.generate(genomes, P, seed=SEED, dense=True)
    Lc, F = synth.to_reference(genomes, P, off, vals)
    ctx = pf.PfaaiGpu(device)
    rec = {"workload": f"all-vs-all {genomes} x {P}, real-like (dense) synthetic family, |F| = {int(off[-1])}", "reps": reps}
    try:
        ctx.upload_index(P, genomes, F, Lc)
        ctx.build()
        for name, path in (("lists", pf.PATH_LISTS), ("auto", pf.PATH_AUTO)):
            ctx.set_path(path, 0)
            ms = []
            for _ in range(reps + 1):
                ctx.compute_device(pf.ALL_VS_ALL)
                ms.append(ctx.timings()["tile_ms"])
            tm = ctx.timings()
            r = {"tile_ms": float(min(ms[1:])), "prep_or_pack_ms": float(tm["prep_ms"] + tm["pack_ms"] + tm["dict_ms"]),
                 "kernel": "aji_mma_kernel" if tm["path"] == pf.PATH_MMA else "aji_tile_kernel"}
            if tm["path"] == pf.PATH_MMA:
                r.update({"dense_tau": tm["dense_tau"], "dense_k_sum": tm["dense_k_sum"], "sparse_items": tm["sparse_items"],
                          "macs_per_s": float(tm["n_tiles"]) * 128 * 128 * tm["dense_k_sum"] / (r["tile_ms"] / 1e3)})
            rec[name] = r
        M = ctx.download_matrix()                                    # the picked path's matrix (last compute)
        v = verify_against_oracle(genomes, P, True, verify_rectangles(genomes, 0, 1), lambda rows, cols: M[np.ix_(rows, cols)], True)
        v.pop("hit_rate", None)
        rec["verify"] = v
        rec["symmetric_zero_diagonal"] = bool(np.array_equal(M, M.T) and not np.any(np.diag(M)))
        rec["speedup_picked_over_lists"] = rec["lists"]["tile_ms"] / rec["auto"]["tile_ms"]
    finally:
        ctx.close()
    rec["wall_s"] = time.perf_counter() - t0
    return rec


def reference_baseline(P, n_genomes, threads, tmpdir):
    db = make_ref_db(n_genomes, P, tmpdir)
    phases, wall = run_ref_once(db, threads, tmpdir)
    pairs = n_genomes * (n_genomes - 1) // 2
    secs = ref_path_seconds(phases)
    return {"value": pairs / secs, "unit": UNIT, "cores": threads, "kind": "reference",
            "sample": f"unmodified reference par_fastaai.x (oracle/_ref) on synthetic FastAAI SQLite DB {n_genomes} genomes x {P} SCPs "
                      f"all-vs-all ({pairs} pairs), OMP_NUM_THREADS={threads}; path time = its own timers for E construction + sort + JAC/AJI "
                      f"{secs:.2f} s (from host arrays, like the GPU arm; its SQLite ingest took another {ref_ingest_seconds(phases):.2f} s; "
                      f"process wall {wall:.2f} s); phases ms: " + ", ".join(f"{k}={v:.0f}" for k, v in phases.items()),
            "pairs_per_s_wall": pairs / wall}


def run_reference_arm(args, rank, emit):
    if rank != 0:
        return 0
    P = args.scps
    threads = os.cpu_count() or 1
    line = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32+f64",
            "data": "synthetic"}
    with tempfile.TemporaryDirectory(prefix="pfaai_ref_") as tmp:
        n = args.ref_genomes
        if ref_binary() is None:
            # the reference did not compile / travel: time the oracle port instead
            from parfastaai_b200 import synth
            off, vals = synth.generate(2048, P, seed=SEED)
            T = synth.t_matrix(2048, P, off)
            for _ in range(args.warmup):
                port_baseline(P, off, vals, T, 512, 512)
            vs = [port_baseline(P, off, vals, T, 2048, 2048) for _ in range(args.steps)]
            v = float(np.mean([x["value"] for x in vs]))
            cb = dict(vs[-1], value=v)
            line.update(value=v, ms_per_step=None, cpu_baseline=cb,
                        config={"workload": "oracle port on synthetic 2048 genomes x %d SCPs" % P})
        else:
            db = make_ref_db(n, P, tmp)
            budget = 240.0
            phases, wall = run_ref_once(db, threads, tmp)          # calibration (also the first warm-up)
            if wall * (args.steps + args.warmup) > budget and n > 24:
                n2 = max(24, int(n * (budget / (wall * (args.steps + args.warmup))) ** 0.5))
                log(f"[reference arm] {n} genomes takes {wall:.1f} s/step; shrinking the sample to {n2} genomes")
                n = n2
                db = make_ref_db(n, P, tmp)
                run_ref_once(db, threads, tmp)
            for _ in range(max(0, args.warmup - 1)):
                run_ref_once(db, threads, tmp)
            pairs = n * (n - 1) // 2
            secs, walls, last = [], [], None
            for _ in range(args.steps):
                last, w = run_ref_once(db, threads, tmp)
                secs.append(ref_path_seconds(last)); walls.append(w)
            v = pairs * len(secs) / sum(secs)
            cb = {"value": v, "unit": UNIT, "cores": threads, "kind": "reference",
                  "sample": f"unmodified reference par_fastaai.x (oracle/_ref), synthetic FastAAI SQLite DB {n} genomes x {P} SCPs all-vs-all "
                            f"({pairs} pairs/step), OMP_NUM_THREADS={threads}; step time = its own timers for E construction + sort + JAC/AJI (from host arrays, like the "
                            f"GPU arm; SQLite ingest not counted); "
                            f"last step phases ms: " + ", ".join(f"{k}={x:.0f}" for k, x in last.items()),
                  "pairs_per_s_wall": pairs * len(walls) / sum(walls)}
            line.update(value=v, ms_per_step=1e3 * sum(secs) / len(secs), cpu_baseline=cb,
                        config={"workload": f"synthetic {n} genomes x {P} SCPs all-vs-all (bounded sample of the bench workload; the "
                                            f"reference cannot run BASELINE configs 4/5: |E| overflows int at ~550 genomes)",
                                "genomes": n, "scps": P})
    line["e2e"] = {"value": line["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    line["gpu_launches"] = 0
    emit(line)
    return 0


# ---------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--genomes", type=int, default=0, help="override the workload size (default: 50000 = BASELINE config 5 at every N)")
    ap.add_argument("--scps", type=int, default=122)
    ap.add_argument("--ref-genomes", type=int, default=64, help="size of the reference arm's bounded sample")
    ap.add_argument("--cpu-baseline-genomes", type=int, default=96)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-dense-family", action="store_true", help="N=1: skip the untimed dense-family side record (both tile paths at config-4 size)")
    ap.add_argument("--no-ceiling", action="store_true", help="skip the on-chip ceiling micro-benchmark (pfaai_gpu_microbench)")
    ap.add_argument("--replicated-build", action="store_true",
                    help="N>1: every rank builds the whole index (the pre-partitioning path) instead of pfaai_gpu_build_partitioned")
    ap.add_argument("--sliced-index", action="store_true",
                    help="N>1: routed build -- every rank keeps only its 1/N slice of F resident, the postings travel to the rank that "
                         "owns their genome by an NCCL all-to-all (no all-gather of F, no filter passes).  Default at N >= 8, where it "
                         "was measured faster end to end (316.5 vs 335.4 ms per step at config 5, same device-resident step); at 2 and 4 "
                         "GPUs the partitioned build on the all-gathered F is at least as fast (DESIGN.md section 6)")
    ap.add_argument("--no-sliced-index", action="store_true", help="N>=8: the partitioned build on the all-gathered F instead of the routed build")
    ap.add_argument("--no-overlap", action="store_true", help="N>1: finish the dictionary exchange before the tile kernel starts")
    ap.add_argument("--verify", action="store_true", help="N>1: gather the shards on rank 0 and compare with an unsharded compute")
    ap.add_argument("--dense", action="store_true",
                    help="the dense (real-like) synthetic family of SURVEY 8(d): 4 clades, d_k ~ U(0, 0.05), e_g ~ U(0, 0.01)")
    ap.add_argument("--no-shared-matrix", action="store_true",
                    help="N>1: keep compact per-rank matrices (upper-triangle cells only) instead of one shared host matrix")
    ap.add_argument("--tau", type=int, default=0, help="dense path: posting-list length threshold (0 = cost model)")
    ap.add_argument("--path", default="auto", choices=["auto", "lists", "mma"],
                    help="tile path: lists = dictionary walk (K4), mma = dense tcgen05 contraction + in-kernel sparse walk (K5), auto = by density")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        log("[bench] note: fewer than 3 warm-up steps requested; the timing rules ask for W >= 3")

    # stdout carries exactly ONE JSON line: everything else that lands on fd 1 (e.g. NCCL's version banner) goes to stderr
    json_fd = os.dup(1)
    os.dup2(2, 1)

    def emit(obj):
        os.write(json_fd, (json.dumps(obj) + "\n").encode())

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference_arm(args, rank, emit)

    import torch
    import torch.distributed as dist

    import __graft_entry__ as ge
    if rank == 0 and not os.path.exists(ge.LIB):
        ge.build()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
        dist.barrier()
    import parfastaai_b200 as pf
    from parfastaai_b200 import sharding, synth

    P = args.scps
    G = args.genomes or 50000
    pairs_total = G * (G - 1) // 2

    class _DevMem:   # raw device pointer of the C-ABI context -> torch tensor (for copies / NCCL on the index buffers)
        def __init__(self, ptr, nbytes):
            self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}

    # ---- synthetic workload (reference-format arrays Lc, F) in pinned host memory; only rank 0 holds it ----
    t0 = time.perf_counter()
    off = vals = None
    nF_t = torch.zeros(1, dtype=torch.int64, device=dev)
    if rank == 0:
        if world > 1:
            # torchrun exports OMP_NUM_THREADS=1; the synthetic generator (host tooling, OpenMP) should use the box
            try:
                import ctypes
                ctypes.CDLL("libgomp.so.1").omp_set_num_threads(max(1, (os.cpu_count() or 1)))
            except OSError:
                pass
        off, vals = synth.generate(G, P, seed=SEED, dense=args.dense)
        nF = int(off[-1])
        F_pin = torch.empty(nF * 2, dtype=torch.int32).pin_memory()
        Lc_pin = torch.empty(pf.NTETRA, dtype=torch.int32).pin_memory()
        F = F_pin.numpy().view(pf.PG_DTYPE)
        Lc = Lc_pin.numpy()
        synth.to_reference(G, P, off, vals, out_F=F, out_Lc=Lc)
        nF_t[0] = nF
        log(f"[rank 0] synthetic {G} x {P}: |F| = {nF}, generated in {time.perf_counter()-t0:.1f} s")
        if world > 1:
            vals = None   # 7 GB at config 5; not needed any more
    if world > 1:
        dist.broadcast(nF_t, src=0)
    nF = int(nF_t.item())

    stream = torch.cuda.Stream(device=dev)
    ctx = pf.PfaaiGpu(local_rank)
    ctx.set_stream(stream.cuda_stream)
    ctx.set_shard(rank, world)
    ctx.set_path({"auto": pf.PATH_AUTO, "lists": pf.PATH_LISTS, "mma": pf.PATH_MMA}[args.path], args.tau)
    my_rows, my_cols = sharding.shard_shape(pf.ALL_VS_ALL, G, 0, rank, world)
    my_pairs = sharding.pair_count(pf.ALL_VS_ALL, G, 0, rank, world)
    # N > 1: ONE G x G host matrix shared by all ranks (POSIX shared memory, page-locked in every process with
    # cudaHostRegister): every rank's tile kernel stores its cells AND their mirror images straight into it
    # (pfaai_gpu_set_global_matrix), so the e2e step ends with the complete symmetric matrix in host memory exactly as at N = 1
    # -- equal work, equal bytes (8*G*G) at every N, no gather and no host mirror pass.
    shared_matrix, shm_path, shm_note = False, None, None
    if world > 1 and not args.no_shared_matrix:
        shm_path = "/dev/shm/pfaai_bench_M_%s.bin" % os.environ.get("MASTER_PORT", "0")
        ok = torch.ones(1, dtype=torch.int32, device=dev)
        need = (8 * G * G + 4095) & ~4095            # whole pages: the mapping is page-locked page by page
        try:
            if rank == 0:
                st = os.statvfs("/dev/shm")
                if st.f_bavail * st.f_frsize < need + (1 << 30):
                    raise RuntimeError("/dev/shm has %.1f GB free, the matrix needs %.1f GB" % (st.f_bavail * st.f_frsize / 1e9, need / 1e9))
                with open(shm_path, "wb") as fh:
                    fh.truncate(need)
        except Exception as e:   # noqa: BLE001
            shm_note = str(e)
            ok[0] = 0
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 1:
            M_sh = torch.from_file(shm_path, shared=True, size=need // 8, dtype=torch.float64)
            if rank == 0:
                # fault every page in once, before anybody page-locks the mapping (tools/probe_host_register.py: registering a
                # never-touched 20 GB shared mapping from two processes at once failed on this box; a populated one does not)
                M_sh.view(torch.uint8)[::4096] = 0
            dist.barrier()
            t_reg = time.perf_counter()
            ok[0] = 1 if pf.host_register(M_sh.data_ptr(), need) else 0     # 1 GiB pieces; clears the CUDA error on failure
            mine_ok = int(ok.item())
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok.item()) == 1:
                shared_matrix = True
                M = M_sh.numpy()[:G * G].reshape(G, G)     # addressed globally only by the e2e steps (set_global_matrix below)
                log(f"[rank {rank}] shared host matrix: {need/1e9:.1f} GB page-locked in {time.perf_counter()-t_reg:.1f} s")
            else:
                shm_note = "page-locking the shared matrix failed on some rank"
                if mine_ok:
                    pf.host_unregister(M_sh.data_ptr(), need)
                del M_sh
                dist.barrier()
                if rank == 0:
                    try:
                        os.remove(shm_path)
                    except OSError:
                        pass
        if not shared_matrix:
            log(f"[rank {rank}] shared host matrix unavailable ({shm_note}); compact per-rank matrices (upper-triangle cells only)")
    if not shared_matrix:
        M_pin = torch.empty(my_rows * my_cols, dtype=torch.float64).pin_memory()
        M = M_pin.numpy().reshape(my_rows, my_cols)
    if world > 1:
        # N > 1: the contexts expose their device index buffers; F is sized for N equal slices (the last one padded) so that the
        # all-gather of the e2e region lands directly in the context's buffer
        per = (nF + world - 1) // world
        per += (-per) % 2                                   # 16-byte aligned slices
        f_ptr0, _, _ = ctx.index_device_buffers(P, G, per * world)
        f_ptr, lc_ptr, _ = ctx.index_device_buffers(P, G, nF)
        assert f_ptr == f_ptr0, "the context reallocated its F buffer"
        F_all = torch.as_tensor(_DevMem(f_ptr, 8 * per * world), device=dev)
        F_dev = F_all[:max(8 * nF, 8)]
        Lc_dev = torch.as_tensor(_DevMem(lc_ptr, 4 * pf.NTETRA), device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if world > 1:
        # one-time, untimed: hand every rank its 1/N slice of F in its OWN pinned host memory (as a one-process-per-GPU loader
        # would hold it): rank 0 uploads F once, broadcasts it, every rank copies its slice back
        lo, hi = min(nF, rank * per), min(nF, (rank + 1) * per)
        with torch.cuda.stream(stream):
            if rank == 0:
                F_dev.copy_(F_pin.view(torch.uint8), non_blocking=True)
                Lc_dev.copy_(Lc_pin.view(torch.uint8), non_blocking=True)
            dist.broadcast(F_dev, src=0)
            dist.broadcast(Lc_dev, src=0)
            F_slice_pin = torch.empty(8 * per, dtype=torch.uint8).pin_memory()
            Lc_pin_r = torch.empty(4 * pf.NTETRA, dtype=torch.uint8).pin_memory()
            F_slice_pin[:8 * (hi - lo)].copy_(F_dev[8 * lo:8 * hi], non_blocking=True)
            Lc_pin_r.copy_(Lc_dev, non_blocking=True)
        stream.synchronize()
        if rank == 0 and not (args.verify and (args.sliced_index or world >= 8)):
            del F_pin, F
        sliced = (args.sliced_index or (world >= 8 and not args.no_sliced_index)) and not args.replicated_build and args.path != "mma"
        if sliced:
            # from here on the context holds only this rank's slice of F (the buffer of the full F is reused)
            fs_ptr, lcs_ptr = ctx.index_slice_buffers(P, G, nF, lo, hi - lo)
            Fs_dev = torch.as_tensor(_DevMem(fs_ptr, max(8 * (hi - lo), 8)), device=dev)
            Lcs_dev = torch.as_tensor(_DevMem(lcs_ptr, 4 * pf.NTETRA), device=dev)
            with torch.cuda.stream(stream):
                Fs_dev[:8 * (hi - lo)].copy_(F_slice_pin[:8 * (hi - lo)], non_blocking=True)
                Lcs_dev.copy_(Lc_pin_r, non_blocking=True)
            stream.synchronize()
    else:
        sliced = False

    def replicate_index():
        """host -> every GPU.  N = 1: pfaai_gpu_upload_index.  N > 1: every rank copies ITS slice of F H2D over its own
        PCIe link, the slices are all-gathered over NVLink / NVSwitch (NCCL) into every context's index buffer."""
        if world == 1:
            ctx.upload_index(P, G, F, Lc)
            return
        cur = torch.cuda.current_stream(dev)               # the copies are torch's: timed here (the C-ABI's h2d_ms sees only its own)
        h_ev[0].record(cur)
        if sliced:                                         # only the rank's own slice, over its own PCIe link
            Fs_dev[:8 * (hi - lo)].copy_(F_slice_pin[:8 * (hi - lo)], non_blocking=True)
            Lcs_dev.copy_(Lc_pin_r, non_blocking=True)
            h_ev[1].record(cur); h_ev[2].record(cur)
            return
        mine = F_all[8 * per * rank:8 * per * (rank + 1)]
        mine.copy_(F_slice_pin, non_blocking=True)
        Lc_dev.copy_(Lc_pin_r, non_blocking=True)
        h_ev[1].record(cur)
        dist.all_gather_into_tensor(F_all, mine)           # in place, straight into the context's F buffer
        h_ev[2].record(cur)

    h_ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]   # N > 1: slice upload begin / end, index all-gather end
    partitioned = world > 1 and not args.replicated_build and args.path != "mma"   # the dense path needs every genome's lists on the GPU
    xviews = {}
    xstream = torch.cuda.Stream(device=dev) if world > 1 else None
    x_ev = [torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)]

    rviews = {}

    def route_and_build():
        """Routed build: group the slice's items by owner, all-to-all over NVLink, sort + own dictionaries on the receiver."""
        send_ptr, offs = ctx.route_prepare()
        with torch.cuda.stream(stream):
            cnt = torch.tensor([offs[d + 1] - offs[d] for d in range(world)], dtype=torch.int64, device=dev)
            allc = torch.empty(world * world, dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(allc, cnt)
            allc = allc.view(world, world).cpu()             # [source, destination]
            send_counts = [int(x) for x in allc[rank]]
            recv_counts = [int(allc[s_, rank]) for s_ in range(world)]
            total = sum(recv_counts)
            recv_ptr = ctx.route_recv_buffer(total)
            for key in ((send_ptr, offs[-1]), (recv_ptr, total)):
                if key not in rviews:
                    rviews[key] = torch.as_tensor(_DevMem(key[0], max(8 * key[1], 8)), device=dev)
            send, recv = rviews[(send_ptr, offs[-1])][:8 * offs[-1]], rviews[(recv_ptr, total)][:8 * total]
            dist.all_to_all_single(recv, send, output_split_sizes=[8 * x for x in recv_counts],
                                   input_split_sizes=[8 * x for x in send_counts])
        ctx.build_routed(total)

    def build_index():
        """N = 1 (or --replicated-build): pfaai_gpu_build.  N > 1: every rank sorts the postings and builds the column
        dictionaries of ITS OWN row blocks only; the four result buffers are all-gathered in place over NVLink (NCCL)."""
        if not partitioned:
            ctx.build()
            return
        if sliced:
            route_and_build()
        else:
            ctx.build_partitioned()              # returns with this shard's chunks complete (host-synchronous)
        # the exchange runs on its own stream; the context is told the event that ends it, runs the tiles on its own dictionaries
        # meanwhile and orders the remaining tiles behind the event (pfaai_gpu_finish_partitioned_async)
        with torch.cuda.stream(xstream):
            x_ev[0].record(xstream)
            # small buffers first: their collectives are over before the tile kernel occupies every SM; the one that moves
            # the dictionary blobs is then already resident and shares the GPU with the tiles on the rank's own dictionaries
            for ptr, chunk, _used in sorted(ctx.exchange_buffers(), key=lambda b: b[1]):
                key = (ptr, chunk)
                if key not in xviews:
                    xviews[key] = torch.as_tensor(_DevMem(ptr, chunk * world), device=dev)
                buf = xviews[key]
                dist.all_gather_into_tensor(buf, buf[chunk * rank:chunk * (rank + 1)])
            x_ev[1].record(xstream)
        if args.no_overlap:
            stream.wait_event(x_ev[1])
            ctx.finish_partitioned()
        else:
            ctx.finish_partitioned_async(x_ev[1].cuda_event)

    def step_device():
        build_index()
        ctx.compute_device(pf.ALL_VS_ALL)

    def step_e2e():
        replicate_index()
        build_index()
        # one host-buffer call; M is pinned, so the tile kernel stores the (shard's) matrix straight into it
        ctx.compute(pf.ALL_VS_ALL, want_jac=False, matrix_out=M)

    def timed(fn, steps, sampler_index=None):
        tile_ms, launches = [], 0
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        clk = ClockSampler(sampler_index) if sampler_index is not None else None
        if clk:
            clk.__enter__()
        with torch.cuda.stream(stream):
            e0.record(stream)
            for _ in range(steps):
                fn()
                tm = ctx.timings()
                tile_ms.append(tm["tile_ms"]); launches += tm["n_launches"]
            e1.record(stream)
        barrier()
        if clk:
            clk.__exit__(None, None, None)
        ms = max_over_ranks(e0.elapsed_time(e1))
        return ms, tile_ms, launches, (clk.summary() if clk else None), tm

    # ---- device-resident metric ----
    if world == 1:
        replicate_index()
    if partitioned:
        # untimed, once: size the shards' dictionary chunks to what this database needs (max over the ranks + 2 %), so that
        # the in-place all-gather of the timed steps moves almost no padding
        with torch.cuda.stream(stream):
            if sliced:
                route_and_build()
            else:
                ctx.build_partitioned()
            used = torch.tensor([float(ctx.exchange_buffers()[0][2])], dtype=torch.float64, device=dev)
            dist.all_reduce(used, op=dist.ReduceOp.MAX)
            ctx.set_partition_chunk(int(used.item() * 1.02) + 4096)
    for _ in range(args.warmup):
        step_device()
    ms_dev, tile_ms, launches, clocks, tm = timed(step_device, args.steps, sampler_index=local_rank)
    exchange_ms = max_over_ranks(x_ev[0].elapsed_time(x_ev[1])) if partitioned else 0.0
    exchange_bytes = sum(ch for _p, ch, _u in ctx.exchange_buffers()) * world if partitioned else 0
    value = pairs_total * args.steps / (ms_dev / 1e3)
    # ---- end to end through the host-buffer C-ABI ----
    if shared_matrix:
        ctx.set_global_matrix(True)
    for _ in range(min(args.warmup, 2)):
        with torch.cuda.stream(stream):
            step_e2e()
    ms_e2e, _, _, _, tm_e2e = timed(step_e2e, args.steps)
    # N > 1: what the e2e step adds in front of the device-resident step, from the LAST e2e step, max over ranks: the rank's own
    # slice of F over its PCIe link (all N links at once) and, for the builds that hold all of F, the NVLink all-gather of the slices
    e2e_h2d_n = max_over_ranks(h_ev[0].elapsed_time(h_ev[1])) if world > 1 else None
    e2e_index_allgather = max_over_ranks(h_ev[1].elapsed_time(h_ev[2])) if world > 1 else None
    e2e_value = pairs_total * args.steps / (ms_e2e / 1e3)
    h2d = 8 * nF + 4 * pf.NTETRA * world              # all ranks together (N > 1: every rank uploads 1/N of F)
    # bytes the tile kernels actually store into host memory per step, all ranks together: N = 1 and the shared matrix hold
    # both triangles + the diagonal (8*G*G); compact per-rank matrices hold the upper-triangle cells + the diagonal only
    if world == 1 or shared_matrix:
        d2h = 8 * G * G
    else:
        d2h = 8 * (pairs_total + G)
    launches = int(max_over_ranks(float(launches)))

    # ---- parity evidence on the path that was timed (every rank, every N): the oracle re-computes rectangles of this rank's
    #      share from re-generated lists and compares them with what the last e2e step stored in host memory ----
    own_ids = sharding.owned_indices(G, rank, world)
    if world == 1 or shared_matrix:
        def read_cells(rows, cols):
            return M[np.ix_(rows, cols)]
    else:
        local_of = {int(g): k for k, g in enumerate(own_ids)}

        def read_cells(rows, cols):
            return M[np.ix_([local_of[int(g)] for g in rows], cols)]
    t0 = time.perf_counter()
    try:
        import ctypes
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(max(1, (os.cpu_count() or 1) // world))
    except OSError:
        pass
    vq = verify_against_oracle(G, P, args.dense, verify_rectangles(G, rank, world), read_cells,
                               lower_valid=(world == 1 or shared_matrix))
    vq["seconds"] = time.perf_counter() - t0
    if world > 1:
        vt = torch.tensor([vq["max_abs_diff"]], dtype=torch.float64, device=dev)
        dist.all_reduce(vt, op=dist.ReduceOp.MAX)
        ct = torch.tensor([float(vq["cells"]), vq["seconds"]], dtype=torch.float64, device=dev)
        cmax = ct.clone()
        dist.all_reduce(ct)
        dist.all_reduce(cmax, op=dist.ReduceOp.MAX)
        gathered = [None] * world
        dist.all_gather_object(gathered, vq["rects"])
        vq = {"max_abs_diff": float(vt.item()), "cells": int(ct[0].item()), "seconds": float(cmax[1].item()),
              "rects": {f"rank{r}": g for r, g in enumerate(gathered)}, "hit_rate": vq.get("hit_rate")}
    vq.update(ranks=world, against="oracle pfo_direct_aji on lists re-generated per rank (seed, genome, SCP)",
              matrix="the host matrix of the last e2e step" + (" (shared by all ranks, mirror images included)" if shared_matrix else ""))
    log(f"[rank {rank}] verify: max |AJI - oracle| = {vq['max_abs_diff']:.3g} over {vq['cells']} cells in {vq['seconds']:.1f} s")

    verify = None
    if args.verify:
        # gather the shards' compact matrices on rank 0, assemble (mirror) and compare with an unsharded compute
        lay = sharding.owned_indices(G, rank, world)
        if world == 1 or shared_matrix:
            full = M
        elif rank == 0:
            parts, lays = [M.copy()], [lay]
            for r in range(1, world):
                shp = sharding.shard_shape(pf.ALL_VS_ALL, G, 0, r, world)
                buf = torch.empty(shp, dtype=torch.float64, device=dev)
                dist.recv(buf, src=r)
                parts.append(buf.cpu().numpy()); lays.append(sharding.owned_indices(G, r, world))
            full = sharding.assemble(pf.ALL_VS_ALL, parts, lays, np.empty((G, G)))
        else:
            dist.send(torch.from_numpy(M).to(dev), dst=0)
        if rank == 0:
            full = np.array(full, copy=True)
            ctx.set_global_matrix(False)
            ctx.set_shard(0, 1)
            if sliced:
                ctx.upload_index(P, G, F, Lc)             # rank 0 kept the whole F for this check
            ctx.build()                                   # the replicated, unsharded path on the same F
            ref = ctx.compute(pf.ALL_VS_ALL, want_jac=False)["matrix"]
            ctx.set_shard(rank, world)
            verify = {"max_abs_diff_vs_unsharded": float(np.abs(full - ref).max()), "identical": bool(np.array_equal(full, ref))}

    # ---- on-chip ceiling of the list path's access pattern, measured on this GPU (rank 0): the walk loop alone against a
    #      shared-memory-resident dictionary at the workload's hit rate ----
    ceiling = None
    if rank == 0 and not args.no_ceiling:
        hr = vq.get("hit_rate") or 0.43
        try:
            c0, c1 = ctx.microbench(0, hr, 30000), ctx.microbench(1, hr, 30000)
            ceiling = {"lookups_per_s": c0["lookups_per_s"], "lookups_per_s_no_bank_conflicts": c1["lookups_per_s"],
                       "hit_rate": hr, "warps_per_sm": c0["warps_per_sm"], "ms": [c0["ms"], c1["ms"]],
                       "what": "pfaai_gpu_microbench: the tile kernel's walk loop (LDS.32 bitmap probe + rank + predicated LDS.64 mask + "
                               "carry-save adds, ids by LDG.128 from L2) on a resident dictionary; no bulk copies, no epilogue, no stores"}
        except Exception as e:   # noqa: BLE001
            ceiling = {"error": str(e)}

    e_size = None
    if world == 1 or not partitioned:
        try:
            e_size = ctx.e_size()          # exact |E| (untimed): the reference's work unit, countTetramerTuples ds_helper.hpp:206-265
        except Exception as e:   # noqa: BLE001
            log(f"[bench] e_size unavailable: {e}")
    peak, peak_src = measured_peaks()
    tile_avg_ms = float(np.mean(tile_ms))
    alg_bytes = int(tm["tile_alg_bytes"])
    achieved = alg_bytes / (tile_avg_ms / 1e3) / 1e9
    traffic = onchip = None
    tp = os.path.join(ROOT, "profiles", f"tile_kernel_traffic_{G}.json")
    if os.path.exists(tp):
        try:
            with open(tp) as fh:
                tj = json.load(fh)
            if tj.get("genomes") == G and tj.get("scps") == P and world == 1:
                traffic = tj.get("dram_bytes_per_launch")
                onchip = tj.get("onchip")
        except Exception:
            traffic = None

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "int32+f64", "data": "synthetic",
        "config": {"workload": f"synthetic {G} genomes x {P} SCPs all-vs-all (BASELINE config {'4' if G == 10000 else '5' if G == 50000 else 'custom'}; "
                               f"SURVEY 8(d) recipe, {'DENSE real-like family, ' if args.dense else ''}seed {SEED})",
                   "family": "dense" if args.dense else "sparse",
                   "genomes": G, "scps": P, "postings": nF, "pairs": pairs_total, "mode": "all-vs-all",
                   "parallelism": (f"output row blocks (384 genomes, folded assignment) sharded over {world} rank(s); "
                                   + ("routed build: every rank uploads and keeps only its 1/N slice of F (inside the e2e region), the postings go "
                                      "to the rank that owns their genome by an NCCL all-to-all; " if sliced else
                                      "F replicated on every GPU (N>1, inside the e2e region: every rank uploads its 1/N slice of F, NCCL all-gather "
                                      "over NVLink); ")
                                   + ("partitioned index build: every rank sorts / builds the column dictionaries of its own row blocks, "
                                      "the dictionaries are all-gathered over NVLink (NCCL, in place, on a second stream) while the tiles on the rank's "
                                      "own dictionaries are computed; " if partitioned else
                                      "index build replicated on every rank; ") +
                                   "no reduction collective; " + ("every rank's tile kernel stores its cells and their mirror images straight into ONE page-locked "
                                                                "host matrix shared by all ranks (POSIX shared memory + cudaHostRegister): the e2e step ends with "
                                                                "the complete symmetric matrix in host memory, as at N = 1" if shared_matrix else
                                                                "every rank stores only its own upper-triangle rows into pinned host memory"
                                                                if world > 1 else "the tile kernel stores both triangles straight into pinned host memory")),
                   "l2_policy": "inputs larger than L2 (F = %.2f GB, matrix = %.2f GB vs 126 MB L2)" % (8 * nF / 1e9, d2h / 1e9),
                   "step": "index build (scan + histogram + radix sort) + row-slab packing + column dictionaries"
                           + (" + dictionary all-gather" if partitioned else "") + " + tile kernel"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / args.steps},
        "gpu_launches": int(launches),
        "clocks": {"sm_mhz": clocks["sm_mhz"], "sm_max_mhz": clocks["sm_max_mhz"], "reasons": clocks["reasons"],
                   "samples": clocks["samples"], "power_w_max": clocks.get("power_w_max")},
        "path": {"tile_kernel": "aji_mma_kernel (K5: tcgen05 u8 dense contraction + in-kernel sparse walk)" if tm["path"] == pf.PATH_MMA
                 else "aji_tile_kernel (K4: dictionary walk, bit-sliced counters)",
                 "requested": args.path, "dense_tau": tm["dense_tau"], "dense_k_sum": tm["dense_k_sum"],
                 "sparse_items": tm["sparse_items"], "prep_ms": tm["prep_ms"]},
        "roofline": {"bound": "hbm", "kernel": "aji_mma_kernel" if tm["path"] == pf.PATH_MMA else "aji_tile_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "alg_bytes_per_launch": alg_bytes, "kernel_ms": tile_avg_ms,
                     "note": "the tile kernel is on-chip (shared-memory lookup / LOP3 issue) bound at this size; see DESIGN.md section 4"},
        # the limiter of the tile kernel is on-chip: live work rates + the pipe utilisations of the committed
        # ncu capture of this same configuration (profiles/tile_kernel_traffic.json)
        "roofline_onchip": {"lookups_per_s": (float(tm["n_tiles"]) * 384 * (nF / G) / (tile_avg_ms / 1e3)) if tm["path"] == pf.PATH_LISTS else None,
                            "pair_scp_updates_per_s": (my_pairs if world > 1 else pairs_total) * P / (tile_avg_ms / 1e3),
                            "lookup": "one row tetramer against one 64-column dictionary (LDS.32 + rank + LDS.64 + bit-sliced add)",
                            "e_size": e_size,
                            "e_per_s": (e_size / world / (tile_avg_ms / 1e3)) if e_size else None,
                            "e_unit": "|E| = sum over (SCP, tetramer) posting lists of n(n-1)/2 pair increments (the reference's E array); per GPU",
                            "ceiling": ceiling,
                            "frac": ((float(tm["n_tiles"]) * 384 * (nF / G) / (tile_avg_ms / 1e3)) / ceiling["lookups_per_s"]
                                     if ceiling and ceiling.get("lookups_per_s") and tm["path"] == pf.PATH_LISTS else None),
                            "frac_of": "the walk loop's own ceiling (the same loop on the same data structure, no bulk copies / epilogue / output); "
                                       "it bounds the overhead around the walk -- it is NOT a roofline fraction (that is roofline.frac)",
                            "mma": ({"macs_per_s": float(tm["n_tiles"]) * 128 * 128 * tm["dense_k_sum"] / (tile_avg_ms / 1e3),
                                     "operand_bytes_per_s": float(tm["n_tiles"]) * 256 * tm["dense_k_sum"] / (tile_avg_ms / 1e3),
                                     "note": "u8 x u8 -> s32 multiply-accumulates on tcgen05 (0/1 operands: exact counts) and the bulk-copied operand bytes they consume"}
                                    if tm["path"] == pf.PATH_MMA else None),
                            "ncu": onchip},
        "phases_ms": {"build": tm["build_ms"], "sort": tm["sort_ms"], "plan": tm["plan_ms"], "pack": tm["pack_ms"], "dict": tm["dict_ms"], "tile": tile_avg_ms,
                      "exchange": exchange_ms, "exchange_bytes_per_rank": exchange_bytes,
                      "e2e_h2d": tm_e2e["h2d_ms"] if world == 1 else e2e_h2d_n, "e2e_d2h": tm_e2e["d2h_ms"],
                      "e2e_h2d_gbs_per_rank": (8.0 * (nF / world) / (e2e_h2d_n / 1e3) / 1e9) if world > 1 and e2e_h2d_n else None,
                      "e2e_index_allgather": e2e_index_allgather,
                      # the same tile kernel inside the e2e step, where its stores go to page-locked HOST memory (rank 0's last step):
                      # the difference to "tile" is what the host-side write bandwidth costs when N GPUs store 8*G*G bytes at once
                      "e2e_tile": tm_e2e["tile_ms"], "e2e_build": tm_e2e["build_ms"]},
    }

    line["verify"] = vq
    if verify is not None:
        line["verify_unsharded"] = verify
    # ---- CPU baseline (rank 0, N=1 only): the reference itself on a bounded sample, plus our OpenMP port ----
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        T = synth.t_matrix(G, P, off)
        try:
            port = port_baseline(P, off, vals, T, 1024, 2048, gpu_matrix=M)
        except Exception as e:  # the checker must not take the bench down
            port = {"error": str(e)}
        cb = None
        if ref_binary() is not None:
            try:
                with tempfile.TemporaryDirectory(prefix="pfaai_ref_") as tmp:
                    cb = reference_baseline(P, args.cpu_baseline_genomes, os.cpu_count() or 1, tmp)
            except Exception as e:
                log(f"[bench] reference binary unusable here ({e}); reporting the oracle port as cpu_baseline")
        line["cpu_baseline"] = cb if cb is not None else port
        line["cpu_baseline_port"] = port
        zero_share = int(np.count_nonzero(M == 0.0)) - G
        line["config"]["zero_share_pairs"] = zero_share // 2

    ctx.close()
    if rank == 0 and world == 1 and not args.dense and not args.no_dense_family:
        try:
            del M
            line["dense_family"] = dense_family_record(pf, synth, local_rank, P)
        except Exception as e:  # noqa: BLE001  (a side record must not take the bench line down)
            line["dense_family"] = {"error": str(e)}
    if world > 1:
        dist.barrier()
        if shared_matrix:
            pf.host_unregister(M_sh.data_ptr(), need)
            if rank == 0:
                try:
                    os.remove(shm_path)
                except OSError:
                    pass
        dist.destroy_process_group()
    if rank == 0:
        emit(line)
    return 0


if __name__ == "__main__":
    sys.exit(main())
