#!/usr/bin/env python
"""Developer tool: executed warp instructions of k_score2 per stage / helper, from an ncu capture.

    python tools/ncu_budget.py gpurun_out/prof_r2d.ncu-rep [videos]

Reads the source page of the report (``--import-source on``) and adds up the per-line counters
over the line ranges of cst_score2.cuh / cst_rows.cuh that make up each stage (found by their anchor text, so
the table follows the source).  Output: share of executed warp instructions, instructions per
video, share of stall samples, active threads per instruction.
"""
import csv
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, 'cst_captioning_b200', 'csrc')

# (category, anchor that starts it) in file order; a category runs to the next anchor of its file
ROW_ANCHORS = [
    ('canonicalise a row (v2_canon)', 'int v2_canon(u32 u'),
    ('prefix scan of row lengths (v2_scan, v2_seg)', 'struct V2Scan {'),
    ('n-gram keys, tf, list stores (v2_emit)', 'void v2_emit(const u16* ctok'),
    ('video-table insert (v2_insert)', 'u32 v2_step(u32 h)'),
    ('video-table find (v2_find)', 'BOUNDED: the table came from'),
]
ANCHORS = [
    ('df-table probe (v2_match4, v2_resolve)', 'int v2_match4(const u64'),
    ('process: slot decode, idf, tf-idf, store', 'void v2_process(const ScoreParams'),
    ('process: presence bits -> table (references)', 'if (EXCL) {'),
    ('process: bit walk (candidates)', 'const int idx = v2_find<T, BOUNDED>(tab, h, key);'),
    ('process: slot store', 'double2* out = reinterpret_cast<double2*>(list + i);'),
    ('prologue, per-video set-up, prefetch', 'template <int MODE, int NW, int CTAS, int ESZ>'),
    ('stage: row loads + canonicalise call', '// ---- load + canonicalise the stage'),
    ('stage: scan + emit loop', '// ---- scan + emit'),
    ('stage: process call + barriers', '// ---- process ----'),
    ('stage: segment sums (norms / numerators)', '// ---- thread per (row, order): its list segment'),
    ('stage bookkeeping, image write (cook)', 'CST2_TICK(12);'),
    ('scores, baseline, rewards, statistics', '// ---- scores, baseline (utils.py'),
]


def main():
    rep = sys.argv[1]
    videos = int(sys.argv[2]) if len(sys.argv) > 2 else 65536
    by_file = {}
    for fname, anchors in (('cst_score2.cuh', ANCHORS), ('cst_rows.cuh', ROW_ANCHORS)):
        lines = open(os.path.join(CSRC, fname)).read().splitlines()
        starts = []
        for name, anchor in anchors:
            idx = next(i for i, l in enumerate(lines) if anchor in l) + 1
            starts.append((idx, name))
        by_file[fname] = sorted(starts)
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'cuda,sass'],
                         capture_output=True, text=True).stdout
    cur, per = None, {}
    for r in csv.reader(raw.splitlines()):
        if len(r) >= 2 and r[0] == 'File Path':
            cur = r[1].split('/')[-1]
        elif len(r) > 8 and r[0].isdigit():
            try:
                per[(cur, int(r[0]))] = (int(r[7]), int(r[4]), int(r[8]))
            except ValueError:
                pass
    tot = sum(v[0] for v in per.values())
    smp = sum(v[1] for v in per.values())
    acc = {}
    for (f, ln), v in per.items():
        if f in by_file:
            cat = 'other'
            for s0, name in by_file[f]:
                if ln >= s0:
                    cat = name
        elif f == 'cst_common.cuh':
            cat = 'hash (cst_hash, cst_slot)' if ln < 62 else 'df-table probe (v2_match4, v2_resolve)' if ln < 120 else \
                'video-table helpers (cst_common.cuh)'
        elif 'atomic' in f:
            cat = 'video-table insert (v2_insert)'
        elif f.startswith('sm_'):
            cat = 'shuffle / ballot / match intrinsics (all stages)'
        else:
            cat = 'fold of the statistics, misc (%s)' % f
        a = acc.setdefault(cat, [0, 0, 0])
        for i in range(3):
            a[i] += v[i]
    print('%-52s %7s %10s %8s %9s' % ('stage / helper', 'inst %', 'inst/video', 'stall %', 'thr/inst'))
    for cat, a in sorted(acc.items(), key=lambda x: -x[1][0]):
        print('%-52s %6.1f%% %10.0f %7.1f%% %9.1f' % (cat, 100.0 * a[0] / tot, a[0] / videos, 100.0 * a[1] / smp,
                                                       a[2] / max(a[0], 1)))
    print('%-52s %6.1f%% %10.0f' % ('total', 100.0, tot / videos))


if __name__ == '__main__':
    main()
