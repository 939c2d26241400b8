"""Executed warp instructions and stall samples per CUDA source line of one kernel (report captured with
--import-source on; kernels compiled with -lineinfo).   python tools/ncu_linemix.py rep.ncu-rep kernel-regex [top] [smp]   (smp: rank by stall samples)
"""
import collections, csv, io, subprocess, sys


def main():
    rep, rx = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    by_smp = len(sys.argv) > 4 and sys.argv[4] == "smp"
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name",
                          "regex:" + rx], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    fname, hdr, seen_fn = "?", None, None
    cnt, smp, src = collections.Counter(), collections.Counter(), {}
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            if seen_fn is None:
                seen_fn = r[1]
                print("#", r[1][:150])
            elif r[1] != seen_fn:
                break
            continue
        if r[0] == "Line No":
            hdr = r
            iE, iM = hdr.index("Instructions Executed"), hdr.index("# Samples")
            continue
        if hdr is None or len(r) <= iE or r[2] != "-":  # per-line summary rows carry "-" in the SASS address column
            continue
        try:
            key = (fname, int(r[0]))
            cnt[key] += int(r[iE])
            smp[key] += int(r[iM])
            src[key] = r[1].strip()
        except ValueError:
            pass
    tot, ts = sum(cnt.values()), sum(smp.values())
    print(f"# {tot} warp instructions, {ts} stall samples")
    order = [(k, cnt[k]) for k, _ in smp.most_common(top)] if by_smp else cnt.most_common(top)
    for key, n in order:
        print(f"{key[0]:>14s}:{key[1]:<5d} {n:10d} {100 * n / max(tot, 1):5.1f} %  smp {100 * smp[key] / max(ts, 1):5.1f} %  {src[key][:110]}")


main()
