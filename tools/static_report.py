#!/usr/bin/env python
"""Static (no GPU) evidence about the built kernels, written to profiles/r2_static_report.md:

  * ptxas resource usage per kernel (registers, spill bytes, static shared memory) from `nvcc -Xptxas -v` of every .cu,
  * SASS mnemonic counts of libnsb200.so from `cuobjdump -sass` (asynchronous copies LDGSTS, shared/global 128-bit accesses,
    acquire/release accesses of the peer-memory protocol, fp64 FMA), per kernel family.

    python tools/static_report.py

Round 2 closed without an ncu capture; these are the facts that can be read off the binary the benches ran.
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "navier-stokes-solver_b200")
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libnsb200.so")
OUT = os.path.join(ROOT, "profiles", "r2_static_report.md")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def demangle(names):
    res = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True)
    return res.stdout.splitlines() if res.returncode == 0 else list(names)


def short(name, width=150):
    name = re.sub(r"\bnsb::", "", name)
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\((?:[^()]|\([^()]*\))*\)$", "", name)     # drop the argument list
    return name if len(name) <= width else name[:width - 3] + "..."


def ptxas_log(src):
    """`nvcc -Xptxas -v` output of one translation unit (minutes for the SpMV-heavy ones), cached under csrc/_obj by source digest."""
    import hashlib
    h = hashlib.sha1()
    for f in sorted(os.listdir(CSRC)):
        if f == src or f.endswith((".cuh", ".h")):
            h.update(open(os.path.join(CSRC, f), "rb").read())
    cache = os.path.join(CSRC, "_obj", src + ".ptxas." + h.hexdigest()[:12] + ".log")
    if os.path.exists(cache):
        return open(cache).read()
    res = subprocess.run([NVCC] + FLAGS + ["-c", os.path.join(CSRC, src), "-o", "/dev/null"], capture_output=True, text=True)
    if res.returncode != 0:
        raise SystemExit("nvcc failed on %s:\n%s" % (src, res.stderr[-2000:]))
    os.makedirs(os.path.dirname(cache), exist_ok=True)
    with open(cache, "w") as f:
        f.write(res.stderr)
    return res.stderr


def ptxas_rows():
    rows = []
    import concurrent.futures
    srcs = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    with concurrent.futures.ThreadPoolExecutor(max_workers=8) as ex:
        logs = list(ex.map(ptxas_log, srcs))
    for src, log in zip(srcs, logs):
        cur = None
        for ln in log.splitlines():
            m = re.search(r"Compiling entry function '([^']+)' for 'sm_100a'", ln)
            if m:
                cur = {"src": src, "name": m.group(1), "spill_st": 0, "spill_ld": 0, "regs": None, "smem": 0, "stack": 0}
                rows.append(cur)
                continue
            if cur is None:
                continue
            m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", ln)
            if m:
                cur["stack"], cur["spill_st"], cur["spill_ld"] = int(m.group(1)), int(m.group(2)), int(m.group(3))
            m = re.search(r"Used (\d+) registers", ln)
            if m:
                cur["regs"] = int(m.group(1))
                s = re.search(r"(\d+) bytes smem", ln)
                cur["smem"] = int(s.group(1)) if s else 0
    names = demangle([r["name"] for r in rows])
    for r, n in zip(rows, names):
        r["pretty"] = short(n)
    return rows


def sass_counts():
    res = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True)
    if res.returncode != 0:
        raise SystemExit("cuobjdump failed: " + res.stderr[-500:])
    per = collections.OrderedDict()
    cur = None
    pats = collections.OrderedDict([
        ("LDGSTS", r"\bLDGSTS\b"), ("LDG.128", r"\bLDG\.E(?:\.[A-Z0-9_]+)*\.128\b"), ("LDS.128", r"\bLDS(?:\.[A-Z0-9_]+)*\.128\b"),
        ("LDS.64", r"\bLDS(?:\.[A-Z0-9_]+)*\.64\b"), ("DFMA", r"\bDFMA\b"), ("system-scope strong loads (LD*.STRONG.SYS)", r"\bLD[G]?\.E(?:\.[A-Z0-9_]+)*\.STRONG\.SYS\b"),
        ("system-scope fences (MEMBAR.*.SYS)", r"\bMEMBAR\.[A-Z.]*SYS\b"), ("UBLKCP (bulk copy)", r"\bUBLKCP\b"),
        ("UTCMMA/tcgen05", r"\bUTC[A-Z]*MMA\b"), ("STL/LDL (local memory)", r"\b(?:STL|LDL)\b")])
    for ln in res.stdout.splitlines():
        m = re.match(r"\s*Function : (\S+)", ln)
        if m:
            cur = m.group(1)
            per[cur] = collections.Counter()
            continue
        if cur is None or "/*" not in ln:
            continue
        per[cur]["instructions"] += 1 if re.search(r"/\*[0-9a-f]{4}\*/", ln) else 0
        for key, pat in pats.items():
            if re.search(pat, ln):
                per[cur][key] += 1
    return per, list(pats)


def family(pretty):
    for key in ("spmv_sell_kernel", "spmv_csr", "k_bpcg", "k_minres", "k_cg", "k_gs", "k_simple", "k_block", "k_halo", "k_p2p", "k_sell",
                "k_vec", "k_"):
        if key in pretty:
            return key
    return "other"


def main():
    rows = ptxas_rows()
    per, keys = sass_counts()
    with open(OUT, "w") as f:
        f.write("# Static report of the round-2 binary (no GPU involved)\n\n")
        f.write("Generated by `python tools/static_report.py` from the sources at HEAD: `nvcc -Xptxas -v` per translation unit and\n"
                "`cuobjdump -sass` of `libnsb200.so`.  %d kernels (entry functions) in %d translation units.\n\n"
                % (len(rows), len(set(r["src"] for r in rows))))
        spilled = [r for r in rows if r["spill_st"] or r["spill_ld"]]
        f.write("## Spills\n\n")
        if spilled:
            f.write("| kernel | file | registers | spill stores (B) | spill loads (B) |\n|---|---|---|---|---|\n")
            for r in spilled:
                f.write("| `%s` | %s | %d | %d | %d |\n" % (r["pretty"], r["src"], r["regs"], r["spill_st"], r["spill_ld"]))
        else:
            f.write("None: every one of the %d kernels compiles with 0 bytes of spill stores and loads.\n" % len(rows))
        f.write("\n## Register use by kernel family (max / median over the instantiations)\n\n| family | kernels | registers max | registers median "
                "| static smem max (B) |\n|---|---|---|---|---|\n")
        fam = collections.defaultdict(list)
        for r in rows:
            fam[family(r["pretty"])].append(r)
        for k in sorted(fam, key=lambda k: -len(fam[k])):
            regs = sorted(r["regs"] for r in fam[k] if r["regs"] is not None)
            f.write("| `%s*` | %d | %d | %d | %d |\n" % (k, len(fam[k]), regs[-1], regs[len(regs) // 2], max(r["smem"] for r in fam[k])))
        f.write("\nThe SELL-LX SpMV instantiations are launch-bounded to 64 registers (`__launch_bounds__(kSellBlock, NSB_SELL_MINBLOCKS)`), "
                "which is what lets 4 blocks of 256 threads share an SM beside 56 KB of dynamic shared memory each (DESIGN.md 3).\n")
        f.write("\n## The hot-path kernels, one line each\n\n| kernel | file | registers | static smem (B) | stack (B) |\n|---|---|---|---|---|\n")
        hot = [r for r in rows if re.match(r"(k_bpcg|k_minres|k_cg_c|k_gs_|k_simple|k_p2p|k_halo)", r["pretty"])]
        for r in hot:
            f.write("| `%s` | %s | %d | %d | %d |\n" % (r["pretty"], r["src"], r["regs"], r["smem"], r["stack"]))
        f.write("\nSELL-LX SpMV instantiations by fused row epilogue (lanes-per-row T = 1..32 each; `HaloIn` = waits for the halo in-kernel):\n\n"
                "| row op | instantiations | registers max | spills |\n|---|---|---|---|\n")
        byop = collections.defaultdict(list)
        for r in rows:
            m = re.match(r"spmv_sell_kernel<\d+, \d+, (.*), (NoHaloIn|HaloIn)>", r["pretty"])
            if m:
                byop[(m.group(1), m.group(2))].append(r)
        for (op, hin), rs in sorted(byop.items()):
            f.write("| `%s` (%s) | %d | %d | %d |\n" % (op, hin, len(rs), max(r["regs"] for r in rs), sum(r["spill_st"] + r["spill_ld"] for r in rs)))
        tot = collections.Counter()
        famc = collections.defaultdict(collections.Counter)
        names = demangle(list(per))
        for (mangled, cnt), pretty in zip(per.items(), names):
            tot.update(cnt)
            famc[family(short(pretty))].update(cnt)
        f.write("\n## SASS mnemonics in `libnsb200.so` (sm_100a)\n\n| mnemonic | whole library | " +
                " | ".join("`%s*`" % k for k in ("spmv_sell_kernel", "k_bpcg", "k_minres", "k_cg", "k_gs", "k_p2p", "k_halo")) + " |\n|---|---|" +
                "---|" * 7 + "\n")
        for key in ["instructions"] + keys:
            f.write("| %s | %d | " % (key, tot[key]) +
                    " | ".join(str(famc[k][key]) for k in ("spmv_sell_kernel", "k_bpcg", "k_minres", "k_cg", "k_gs", "k_p2p", "k_halo")) + " |\n")
        f.write("\nReading: `LDGSTS` is `cp.async` (global to shared without registers: the staged `x` window and the next chunk's "
                "stream of the SELL-LX kernel); there is no bulk-tensor copy (`UBLKCP`) and no tensor-core instruction, as DESIGN.md 3 "
                "says (fp64 SpMV: HBM-bound, FMA pipe); `LD*.STRONG.SYS` counts every system-scope strong load: the `ld.acquire.sys` polls of the peer-memory halo / all-reduce protocol "
                "(`p2p.cuh`), and also the `__ldcv` / volatile reads of landing buffers, stop flags and the `grid_sum` partials; `MEMBAR.*.SYS` are "
                "the release fences before a flag is published.  Both appear in the SpMV instantiations that wait for a halo in-kernel or push "
                "from their epilogue and in the kernels that finish a reduction in their last block.\n")
    print("wrote", OUT)


if __name__ == "__main__":
    sys.exit(main())
