#!/usr/bin/env python3
"""Joins the SASS page of an ncu report (ncu -i X.ncu-rep --page source --csv) with the line table of the library
(nvdisasm -g) and prints warp instructions executed / stall samples per CUDA source line of one kernel.
usage: ncu_lines.py report.ncu-rep kernel_mangled_substring [top]"""
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "struspattern_b200", "libstrus_pattern_b200.so")


def line_table(mangled):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
    for f in sorted(os.listdir(tmp)):
        sass = subprocess.run(["nvdisasm", "-g", "-c", f], cwd=tmp, stdout=subprocess.PIPE, text=True).stdout
        if mangled not in sass:
            continue
        table, cur, inside = {}, None, False
        for l in sass.split("\n"):
            if l.startswith(".text."):
                inside = mangled in l
                continue
            if not inside:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
            if m:
                inl = re.findall(r'inlined at "([^"]+)", line (\d+)', m.group(3))
                cur = (os.path.basename(m.group(1)), int(m.group(2)), tuple((os.path.basename(a), int(b)) for a, b in inl))
                continue
            m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
            if m:
                table[int(m.group(1), 16)] = (cur, m.group(2))
        if table:
            return table
    raise SystemExit("kernel not found in " + LIB)


def main():
    rep, mangled = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    table = line_table(mangled)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    want = re.sub(r"^_ZN\d*", "", mangled)
    agg, total, hdr, inkernel, base = {}, 0.0, None, False, None
    for r in rows:
        if len(r) >= 2 and r[0] == "Kernel Name":
            inkernel = demangled_matches(r[1], mangled)
            hdr, base = None, None
            continue
        if not inkernel:
            continue
        if r and r[0] == "Address":
            hdr = r
            continue
        if hdr is None or len(r) != len(hdr):
            continue
        d = dict(zip(hdr, r))
        addr = int(d["Address"], 16)
        if base is None:
            base = addr
        key, _ = table.get(addr - base, (None, ""))
        inst = float(d["Instructions Executed"] or 0)
        thr = float(d["Thread Instructions Executed"] or 0)
        smp = float(d["# Samples"] or 0)
        a = agg.setdefault(key, [0.0, 0.0, 0.0, {}])
        a[0] += inst; a[1] += thr; a[2] += smp
        total += inst
        for k in hdr:
            if k.startswith("stall_") and "Not Issued" not in k and d[k] not in ("", "0"):
                a[3][k] = a[3].get(k, 0) + float(d[k])
    print("total warp instructions %.4g" % total)
    for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        stalls = ", ".join("%s %d" % (k[6:], v) for k, v in sorted(a[3].items(), key=lambda kv: -kv[1])[:3])
        where = "?" if key is None else "%s:%d%s" % (key[0], key[1], "".join(" <- %s:%d" % x for x in key[2][:2]))
        print("%5.1f%%  inst %.3g  thr/inst %4.1f  samples %6d  %-60s %s" % (100 * a[0] / total, a[0], a[1] / max(a[0], 1), a[2], where, stalls))


def demangled_matches(name, mangled):
    if "k_eval_passILb0" in mangled:
        return "k_eval_pass<(bool)0>" in name
    if "k_eval_passILb1" in mangled:
        return "k_eval_pass<(bool)1>" in name
    core = re.sub(r"^_ZN?\d*", "", mangled)
    m = re.search(r"(k_[a-z_]+)", mangled)
    return bool(m) and m.group(1) in name


if __name__ == "__main__":
    main()
