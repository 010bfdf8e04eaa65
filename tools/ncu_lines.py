"""Attribute the SASS-level samples of an `ncu --set full --import-source on` capture to source lines and
code sections of ebt_kernels.cu:

    python tools/ncu_lines.py gpurun_out/<name>.ncu-rep <kernel name fragment> <out.json>

The ncu source page lists one row per SASS instruction (address, stall samples, instructions executed);
`nvdisasm -g` of the library's own cubin gives the source line of every instruction offset.  The library
must be the build that was profiled.  Sections are line ranges of the functions in ebt_kernels.cu, found by
their definitions, so the table follows the source as it changes."""
import collections
import csv
import io
import json
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "event_based_odomety_b200", "csrc", "ebt_kernels.cu")
LIB = os.path.join(ROOT, "event_based_odomety_b200", "libebt_cuda.so")

# section name -> functions whose bodies belong to it
SECTIONS = collections.OrderedDict([
    ("pixel pass (eval_pixels, bicubic taps, Jacobian sums)", ["cr_weights", "sample_taps", "sample_taps_tile", "eval_pixels", "pixel_service", "sb_serve"]),
    ("warp butterfly + block reductions", ["butterfly_slot", "Butterfly", "block_sum", "warp_sum", "red_total"]),
    ("pixel warps waiting for the next request (lm_worker, barriers)", ["lm_worker", "handshake", "partials_arrive", "partials_wait"]),
    ("hand-off by sequence numbers: leaders posting / polling for completion (sb_post, wait_partials)", ["sb_post", "wait_partials", "ld_volatile_u32"]),
    ("pixel-warp jobs of the multi-stream kernel: integration, normalisation + tile staging, window-size sum (mq_serve)", ["mq_serve", "num_events_term"]),
    ("leader: per-optimisation scalar work of the multi-stream kernel (mq_optimize, mq_update_num_of_events)", ["mq_optimize", "mq_update_num_of_events"]),
    ("multi-stream kernel body: pixel-warp polling loop, work queue, candidate walk, ring, state, records (track_multi_kernel)", ["track_multi_kernel", "mq_load_segment"]),
    ("leader: evaluation hand-off + totals + finalisation (leader_eval, post_request)", ["leader_eval", "post_request", "huber"]),
    ("leader: trust-region step (lm_solve, solve4, plus5, scale_system, rcp/rsqrt/sincos)",
     ["lm_solve", "solve4", "plus5", "scale_system", "rcp_fast", "rsqrt_fast", "sqrt_fast", "sincos_fast", "inv_len",
      "norm5", "hidx", "sel_max", "sel_min"]),
    ("integration + normalisation + post-solve (optimize_feature, integrate_window, update_rect)",
     ["optimize_feature", "integrate_window", "update_rect", "in_rect", "to_i32"]),
    ("window-size update (update_num_of_events)", ["update_num_of_events"]),
    ("candidate walk, ring, state load/store, records (track_batch_kernel body)",
     ["track_batch_kernel", "window_need", "window_grow", "load_state", "store_state", "write_record",
      "rebuild_mask_row", "candidate_bounds"]),
    ("patch products / cost map", ["motion_compensated", "predicted_nabla", "cost_map"]),
])


def function_ranges():
    """(first line, last line, name) of every function/struct body in ebt_kernels.cu (brace matching from its header)."""
    src = open(SRC).read().split("\n")
    out = []
    names = [n for fs in SECTIONS.values() for n in fs]
    for i, l in enumerate(src):
        for n in names:
            if (re.search(r"\b(void|double|int|bool|int4|long long|struct)\s+" + re.escape(n) + r"\b", l)
                    or re.match(re.escape(n) + r"\(", l)) and not l.strip().startswith("//"):
                # header may span lines; body starts at the first '{'
                j, depth, seen = i, 0, False
                while j < len(src):
                    if ";" in src[j] and not seen and "{" not in src[j]:
                        break  # a declaration or a call, not a definition
                    depth += src[j].count("{") - src[j].count("}")
                    seen = seen or "{" in src[j]
                    if seen and depth == 0:
                        out.append((i + 1, j + 1, n))
                        break
                    j += 1
    return out


def line_of_offset(kernel_fragment):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, capture_output=True)
    cubin = next(f for f in os.listdir(tmp) if f.startswith("ebt_kernels.") and f.endswith(".cubin"))
    sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
    table, cur, active = {}, None, False
    for l in sass.split("\n"):
        if l.startswith(".text.") or l.startswith("\t.section"):
            active = kernel_fragment in l and ".text" in l
        if not active:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = int(m.group(2)) if m.group(1).endswith("ebt_kernels.cu") else 0
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+\S", l)
        if m and cur is not None:
            table[int(m.group(1), 16)] = cur
    return table


def main():
    rep, frag, out = sys.argv[1:4]
    csvtxt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(csvtxt)))
    hdr = next(r for r in rows if r and r[0] == "Address")
    body = [r for r in rows[rows.index(hdr) + 1:] if len(r) == len(hdr)]
    iS, iI = hdr.index("# Samples"), hdr.index("Instructions Executed")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    base = int(body[0][0], 16)
    table = line_of_offset(frag)
    ranges = function_ranges()

    def section_of(line):
        best = None
        for a, b, n in ranges:
            if a <= line <= b and (best is None or (b - a) < (best[1] - best[0])):
                best = (a, b, n)
        if best is None:
            return "other (library device code, headers)"
        return next(s for s, fs in SECTIONS.items() if best[2] in fs)

    sec = collections.defaultdict(lambda: {"samples": 0, "instructions": 0, "stalls": collections.Counter()})
    for r in body:
        line = table.get(int(r[0], 16) - base, 0)
        s = sec[section_of(line)]
        s["samples"] += int(r[iS] or 0)
        s["instructions"] += int(r[iI] or 0)
        for i in stall_cols:
            s["stalls"][hdr[i]] += int(r[i] or 0)
    tot_s = sum(v["samples"] for v in sec.values()) or 1
    tot_i = sum(v["instructions"] for v in sec.values()) or 1
    res = {k: {"sample_share": v["samples"] / tot_s, "instruction_share": v["instructions"] / tot_i,
               "warp_instructions": v["instructions"],
               "top_stalls": {n: c / max(v["samples"], 1) for n, c in v["stalls"].most_common(3)}}
           for k, v in sorted(sec.items(), key=lambda kv: -kv[1]["samples"])}
    json.dump(res, open(out, "w"), indent=1)
    for k, v in res.items():
        print(f"{v['sample_share']*100:5.1f} % samples {v['instruction_share']*100:5.1f} % instr  {k}  "
              + ", ".join(f"{n} {c*100:.0f}%" for n, c in v["top_stalls"].items()))


if __name__ == "__main__":
    main()
