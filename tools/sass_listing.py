"""SASS evidence listing for the hand-written kernels: for every requested kernel of a built object file, the instructions that
prove what it is - tcgen05 MMAs (UTC*MMA), TMEM loads (LDTM), TMA tensor / bulk copies (UTMALDG / UTMASTG / UBLKCP / UTMAREDG),
mbarrier traffic (SYNCS), system-scope loads / stores / reductions (the peer-memory flags of the fused collectives) and
multimem instructions, and for the warp-level kernels HMMA / LDSM (ldmatrix) / LDGSTS (cp.async) - each with its SASS address, in program order (a listing, not a count).

    python tools/sass_listing.py build/gemm.cu.o 'gemm_tcgen05_kernelILi256ELb0ELb0ELi1ELb1' [...more name fragments] > profiles/x.txt
"""
import re
import subprocess
import sys

KEEP = re.compile(r"\b(UTC[A-Z]*MMA|UTCBAR|UTCCP|LDTM|STTM|UTMALDG|UTMASTG|UTMAREDG|UTMAPF|UBLKCP|UBLKRED|SYNCS|ERRBAR|MEMBAR|CCTL|"
                  r"HMMA[^ ]*|LDSM[^ ]*|LDGSTS[^ ]*|LDGDEPBAR|DEPBAR[^ ]*|MULTIMEM|LDGMC[^ ]*|STGMC[^ ]*|REDGMC[^ ]*|UCGABAR|BAR\.SYNC|BAR\.ARV|ELECT|REDG?\.E[^ ]*|ATOMG?\.E[^ ]*)\b|\.SYS\b|\.STRONG\.GPU|\.MMIO")


def main():
    obj, frags = sys.argv[1], sys.argv[2:]
    out = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True, check=True).stdout.split("\n")
    cur, keep_fn, total, kept, counts = None, False, 0, 0, {}

    def flush():
        if cur is not None and keep_fn:
            top = ", ".join(f"{k} x{v}" for k, v in sorted(counts.items(), key=lambda kv: -kv[1])[:14])
            print(f"    -- {total} SASS instructions, {kept} listed above; mnemonic totals: {top}\n")

    for line in out:
        m = re.search(r"Function : (\S+)", line)
        if m:
            flush()
            cur, total, kept, counts = m.group(1), 0, 0, {}
            keep_fn = any(f in cur for f in frags)
            if keep_fn:
                dem = subprocess.run(["c++filt", cur], capture_output=True, text=True).stdout.strip()
                print("=" * 120 + f"\n{dem[:400]}\n" + "=" * 120)
            continue
        if not keep_fn:
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
        if not m:
            continue
        total += 1
        ins = m.group(2).strip()
        if KEEP.search(ins):
            kept += 1
            mn = ins.split()[1] if ins.startswith("@") else ins.split()[0]
            mn = mn.split(".")[0] + ("." + ".".join(x for x in mn.split(".")[1:3]) if "." in mn else "")
            counts[mn] = counts.get(mn, 0) + 1
            print(f"  /*{m.group(1)}*/  {ins}")
    flush()


if __name__ == "__main__":
    main()
