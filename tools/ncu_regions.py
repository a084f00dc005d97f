#!/usr/bin/env python
"""Per-region share of warp instructions, active lanes per instruction and stall samples of one k_trace launch, from the source page of an
ncu report captured with --import-source on (kernels built with -lineinfo):  tools/ncu_regions.py <report.ncu-rep> <launch index> [...]
Regions are ranges of kernels.cu lines found by their function names, so the table survives edits."""
import re, subprocess, sys
from pathlib import Path

SRC = Path(__file__).resolve().parent.parent / "heatray_b200" / "csrc" / "kernels.cu"


def regions():
    lines = SRC.read_text().splitlines()
    def find(pat, start=0):
        for i in range(start, len(lines)):
            if re.search(pat, lines[i]):
                return i + 1
        raise SystemExit(f"pattern {pat} not found")
    tri0 = find(r"bool triangleTest\("); tri1 = find(r"^struct TraceCounters")
    beg0 = find(r"void begin\(float3 org"); beg1 = find(r"bool hasNodeWork\(\)")
    node0 = find(r"void nodeStep\("); node1 = find(r"void popIfEmpty\(")
    ts0 = find(r"bool triStep\("); ts1 = find(r"bool finished\(\) const")
    k0 = find(r"__global__ void __launch_bounds__\(128, HR_TRACE_MIN_BLOCKS\) k_trace"); k1 = find(r"^// ---- material shading")
    return [("triangle test (triangleTest + triStep)", [(tri0, tri1), (ts0, ts1)]), ("ray setup (Traversal::begin)", [(beg0, beg1)]),
            ("node step (Traversal::nodeStep + byte helpers)", [(node0, node1), (find(r"uint32_t byteOf\("), find(r"uint32_t byteOf\(") + 1), (find(r"float byteToUnitFloat\(") - 1, find(r"float byteToUnitFloat\(") + 3)]),
            ("kernel body: refill / retire / finish / loop control / pop", [(k0, k1), (node1, ts0)])]


def main():
    rep = sys.argv[1]
    regs = regions()
    for launch in sys.argv[2:]:
        out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-id", f":::{launch}"], capture_output=True, text=True).stdout
        # not csv.reader: ncu does not escape the quotes of inline-asm strings in the Source column; every other field is numeric, so split on "," and
        # re-join whatever a quoted source line was cut into
        rows = []
        width = None
        for line in out.splitlines():
            f = line.strip().strip('"').split('","')
            if f and f[0] == "Line No":
                width = len(f)
            if width and len(f) > width:
                extra = len(f) - width
                f = [f[0], '","'.join(f[1:2 + extra])] + f[2 + extra:]
            rows.append(f)
        hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
        hdr = rows[hdr_i]
        ci, ct, cs = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
        name = next(r[1] for r in rows if r and r[0] == "Function Name")
        agg = {n: [0, 0, 0] for n, _ in regs}; other = [0, 0, 0]
        for r in rows[hdr_i + 1:]:
            if not r or not r[0].isdigit():
                continue                                            # SASS rows repeat what their source line already sums
            ln = int(r[0]); v = [int(r[ci] or 0), int(r[ct] or 0), int(r[cs] or 0)]
            for n, spans in regs:
                if any(a <= ln < b for a, b in spans):
                    for k in range(3): agg[n][k] += v[k]
                    break
            else:
                for k in range(3): other[k] += v[k]
        agg["other (inlined helpers: float3 operators, loads)"] = other
        ti, tt, ts = (sum(a[k] for a in agg.values()) for k in range(3))
        print(f"# launch {launch}: {name.split('(')[0].replace('void hr::<unnamed>::', '')}  {ti / 1e6:.1f} M warp instructions, {tt / max(ti, 1):.1f} active lanes per instruction")
        print("| region (kernels.cu) | share of warp instructions | active lanes / instruction | share of stall samples |\n|---|---|---|---|")
        for n, a in agg.items():
            if a[0]:
                print(f"| {n} | {a[0] / ti * 100:.1f} % | {a[1] / a[0]:.1f} | {a[2] / max(ts, 1) * 100:.1f} % |")
        print()


if __name__ == "__main__":
    main()
