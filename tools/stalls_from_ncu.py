"""profiles/r3_fill_unit_stalls.txt: where the warps of the dominant kernel wait, from the source page of an ncu --set full report.
  ncu -i gpurun_out/r3_level3.ncu-rep --page source --csv --kernel-name regex:fill_unit > /tmp/src.csv
  python tools/stalls_from_ncu.py /tmp/src.csv > profiles/r3_fill_unit_stalls.txt
The LAST launch in the file is taken (level 3 of config 3's enumeration)."""
import csv
import sys


def main(path):
    secs, cur = [], None
    for row in csv.reader(open(path)):
        if not row:
            continue
        if row[0] == "Kernel Name":
            cur = {"name": row[1], "hdr": None, "rows": []}
            secs.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = row
        elif cur is not None:
            cur["rows"].append(row)
    s = secs[-1]
    ix = {n: i for i, n in enumerate(s["hdr"])}
    num = lambda r, n: int(float(r[ix[n]] or 0))
    tot = sum(num(r, "# Samples") for r in s["rows"])
    print(f"# {s['name']}: last of {len(secs)} launches in the report, {tot} warp samples")
    print("# stall reasons (all samples)")
    reasons = [n for n in s["hdr"] if n.startswith("stall_") and "Not Issued" not in n]
    for n, v in sorted(((n, sum(num(r, n) for r in s["rows"])) for n in reasons), key=lambda t: -t[1])[:8]:
        print(f"{n:24s} {v:6d}  {100 * v / tot:5.1f} %")
    print("# instructions with the most samples (a long-scoreboard wait is charged to the first instruction that needs the load)")
    print("samples | long_sb | SASS")
    for r in sorted(s["rows"], key=lambda r: -num(r, "# Samples"))[:14]:
        print(f"{num(r, '# Samples'):6d} | {num(r, 'stall_long_sb'):6d} | {r[ix['Source']].strip()}")
    print("# global loads and stores: requests and sectors (divergence = sectors per request)")
    print("op | size | L1 tag requests | L2 sectors | ideal | SASS")
    for r in s["rows"]:
        if r[ix["Address Space"]].strip() == "Global" and num(r, "L1 Tag Requests Global") > 0:
            print(f"{r[ix['Access Operation']]:5s} | {r[ix['Access Size']]:>4s} | {num(r, 'L1 Tag Requests Global'):9d} | "
                  f"{num(r, 'L2 Theoretical Sectors Global'):9d} | {num(r, 'L2 Theoretical Sectors Global Ideal'):9d} | {r[ix['Source']].strip()[:70]}")


if __name__ == "__main__":
    main(sys.argv[1])
