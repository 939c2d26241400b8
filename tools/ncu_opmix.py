"""Executed-instruction mix of one kernel out of an .ncu-rep captured with --import-source on (source page, SASS view):
warp instructions and stall samples per opcode.   python tools/ncu_opmix.py rep.ncu-rep kernel-regex|skip=N [top]
(the regex sees the function's base name only; skip=N picks the N-th launch of the report)
"""
import collections, csv, io, subprocess, sys


def main():
    rep, rx = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    sel = ["--launch-skip", rx[5:], "--launch-count", "1"] if rx.startswith("skip=") else ["--kernel-name", "regex:" + rx]
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + sel, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = None
    cnt, smp, tot, nk = collections.Counter(), collections.Counter(), 0, 0
    for r in rows:
        if r and r[0] == "Kernel Name":
            nk += 1
            if nk > 1:
                break  # first matching launch only
            print("#", r[1][:150])
            continue
        if r and r[0] == "Address":
            hdr = r
            iS, iE, iM = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
            continue
        if hdr is None or len(r) <= iE:
            continue
        op = r[iS].strip().split()
        if not op:
            continue
        o = op[1] if op[0].startswith("@") else op[0]
        parts = o.rstrip(";").split(".")
        key = parts[0] + ("." + parts[1] if len(parts) > 1 and parts[0] in ("LDS", "STS", "LDG", "STG", "MUFU", "LDGSTS", "BAR", "RED", "ATOMS", "ATOMG") else "")
        n = int(r[iE])
        cnt[key] += n
        tot += n
        smp[key] += int(r[iM])
    ts = sum(smp.values())
    print(f"# {tot} warp instructions, {ts} stall samples")
    for o, n in cnt.most_common(top):
        print(f"{o:14s} {n:11d} {100 * n / tot:5.1f} %   samples {100 * smp[o] / max(ts, 1):5.1f} %")


main()
