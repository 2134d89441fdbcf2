"""Per-source-line instruction and stall totals of one kernel in an .ncu-rep (run here, no GPU).

    python tools/ncu_lines.py gpurun_out/x.ncu-rep|x_src_<kernel>.csv k_trace_stream [launch-index] [top-n]

ncu's CSV source page is per SASS instruction without line numbers; nvdisasm -g of the cubin inside
liborb200.so has the same instruction sequence with //## File/line markers.  The two are joined by
instruction offset and summed per (file, line).  The .so must be the build that was profiled."""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.environ.get("ORB_LIB", os.path.join(ROOT, "optix_renderer_b200", "liborb200.so"))  # the build that was profiled


def sass_lines(kernel):
    """offset -> (file, line, text) for the first function whose mangled name contains `kernel`."""
    with tempfile.TemporaryDirectory() as t:
        subprocess.check_call(["cuobjdump", "-xelf", "all", LIB], cwd=t, stdout=subprocess.DEVNULL)
        cubin = [f for f in os.listdir(t) if f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(t, cubin)], capture_output=True, text=True).stdout
    out, cur, inside = {}, ("?", 0), False
    for ln in txt.splitlines():
        if ln.startswith(".text."):
            inside = kernel in ln
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            out[int(m.group(1), 16)] = (cur[0], cur[1], m.group(2).strip())
    return out


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    which = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    if rep.endswith(".csv"):  # the source page exported on the GPU box (tools/ncu_full.sh)
        raw = open(rep).read()
    else:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kernel],
                             capture_output=True, text=True).stdout
    # several launches: blocks start with a "Kernel Name" row
    blocks, cur = [], None
    for row in csv.reader(io.StringIO(raw)):
        if row and row[0] == "Kernel Name":
            cur = []
            blocks.append(cur)
        elif cur is not None:
            cur.append(row)
    rows = blocks[which]
    hdr = rows[0]
    ix = {h: i for i, h in enumerate(hdr)}
    lines = sass_lines(os.environ.get("MANGLED", kernel))
    base = int(rows[1][ix["Address"]], 16)
    agg = collections.defaultdict(lambda: [0, 0, 0, collections.Counter()])
    tot_inst = tot_thr = tot_samp = 0
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    for r in rows[1:]:
        off = int(r[ix["Address"]], 16) - base
        f, l, _ = lines.get(off, ("?", 0, ""))
        inst, thr, samp = int(r[ix["Instructions Executed"]]), int(r[ix["Thread Instructions Executed"]]), int(r[ix["# Samples"]])
        a = agg[(f, l)]
        a[0] += inst
        a[1] += thr
        a[2] += samp
        for s in stall_cols:
            v = int(r[ix[s]] or 0)
            if v:
                a[3][s] += v
        tot_inst += inst
        tot_thr += thr
        tot_samp += samp
    print("kernel %s launch %d: %d warp-inst, %.2f threads/inst, %d samples" % (kernel, which, tot_inst, tot_thr / max(1, tot_inst), tot_samp))
    print("%-18s %6s %7s %6s %7s  %s" % ("file:line", "inst%", "thr/in", "samp%", "", "top stalls"))
    for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
        st = ", ".join("%s %d" % (k[6:], v) for k, v in a[3].most_common(3))
        print("%-18s %6.2f %7.1f %6.2f %7s  %s" % ("%s:%d" % (f, l), 100.0 * a[0] / tot_inst, a[1] / max(1, a[0]), 100.0 * a[2] / max(1, tot_samp), "", st))


if __name__ == "__main__":
    main()
