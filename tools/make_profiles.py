#!/usr/bin/env python
"""Turns the raw outputs of a GPU run (gpurun_out/: bench lines, tools/bench_configs.py lines, the ncu launch list) into the
tracked summaries under profiles/.  Usage: make_profiles.py <round tag, e.g. r02> <prefix of the run's files, e.g. r2n>"""
import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")


def last_json(path):
    with open(path) as f:
        lines = [l for l in f.read().splitlines() if l.startswith("{")]
    return json.loads(lines[-1])


def bench_lines(tag, pre):
    for suffix, name in (("bench", "bench"), ("bench_weak", "bench_weak512"), ("bench_cfg2", "bench_cfg2"), ("bench_cfg3", "bench_cfg3"), ("bench_cfg4", "bench_cfg4"),
                         ("bench_ref", "bench_reference_arm")):
        src = os.path.join(OUT, f"{pre}_{suffix}.json")
        if os.path.exists(src):
            with open(os.path.join(PROF, f"{tag}_{name}.json"), "w") as f:
                f.write(json.dumps(last_json(src)) + "\n")
            print("wrote", f"{tag}_{name}.json")


def launches(tag, pre):
    src = os.path.join(OUT, f"{pre}_launches.csv")
    if not os.path.exists(src):
        return
    rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"'))]
    hdr, rows = rows[0], rows[1:]
    iN, iG, iV = hdr.index("Kernel Name"), hdr.index("Grid Size"), hdr.index("Metric Value")
    groups = collections.OrderedDict()
    first_of = {}
    for r in rows:
        name = r[iN].replace("void ", "").split("(")[0]
        us = float(r[iV]) / 1e3
        first_of.setdefault(name, us)              # the first launch of each kernel belongs to a full device-resident batch
        kind = "full batch" if us >= 0.6 * first_of[name] else "piece of a host-buffer call"
        key = (name, r[iG] + " " + kind if name.split("<")[0] in ("icr::prep_kernel", "icr::estimate_kernel", "icr::narrow_kernel") else r[iG])
        groups.setdefault(key, []).append(us)
    ours = {k: v for k, v in groups.items() if k[0].startswith("icr::")}
    with open(os.path.join(PROF, f"{tag}_launches.md"), "w") as f:
        f.write(f"# Round {tag[1:].lstrip('0')}: ncu launch list of the bench command\n\n"
                "Command (B200): `ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv python bench.py --steps 2 --warmup 3 "
                "--no-cpu-baseline` (after the same command exited 0 without ncu).  Times are per launch, cold-cache and serialised: "
                "the kernel's SHARE of a step is what must agree with bench.py, not the absolute.  Launches are grouped by kernel and "
                "grid size: the 4096-image batches of the device-resident steps, the 1024-image pieces the host-buffer entries cut a "
                "batch into (f32 and u8 entry), and the C++ layer's call.\n\n| kernel | grid | launches | mean us | min us | max us |\n|---|---|---|---|---|---|\n")
        for (name, grid), v in ours.items():
            f.write(f"| `{name}` | {grid} | {len(v)} | {sum(v) / len(v):.1f} | {min(v):.1f} | {max(v):.1f} |\n")
        # share inside a device-resident step: the first prep / estimate / big triple in the list
        first = {}
        for r in rows:
            n = r[iN]
            for k in ("prep_kernel", "estimate_kernel", "narrow_kernel", "big_kernel"):
                if k in n and k not in first:
                    first[k] = float(r[iV]) / 1e3
        tot = sum(first.values())
        f.write("\nShare of one device-resident step (first launch of each kernel): " +
                ", ".join(f"{k} {v:.1f} us ({100 * v / tot:.1f} %)" for k, v in first.items()) + f"; sum {tot:.1f} us.\n")
        others = {k: v for k, v in groups.items() if not k[0].startswith("icr::")}
        f.write("\nOther kernels in the run (torch: fills, copies for the checks): " +
                "; ".join(f"`{k[0][:60]}` x{len(v)} ({sum(v) / len(v):.1f} us)" for k, v in list(others.items())[:8]) + "\n")
    print("wrote", f"{tag}_launches.md")


def configs(tag, pre):
    src = os.path.join(OUT, f"{pre}_configs.jsonl")
    if not os.path.exists(src):
        return
    lines = [json.loads(l) for l in open(src) if l.startswith("{")]
    with open(os.path.join(PROF, f"{tag}_configs.md"), "w") as f:
        f.write(f"# Round {tag[1:].lstrip('0')}: the other configs through the C ABI (inputs resident in HBM)\n\n"
                "Command (B200): `python tools/bench_configs.py 2 3 4 big bg i64` -- CUDA events between the library's launches, 10 steps after "
                "3 warm-ups, spot parity against the CPU oracle on the first maps of every config.  `big` (one whole-image segment per "
                "map), `bg` (BSP rectangles on a border-touching background segment) and `i64` are not BASELINE configs: they put the "
                "big-segment kernel and the int64 label path on a measured line.\n\n"
                "| config | label maps | mean K | prep us | estimator us | big+reward us | images/s | rewards/s | algorithmic GB/s | max rel err vs oracle |\n"
                "|---|---|---|---|---|---|---|---|---|---|\n")
        for d in lines:
            if "kernel_ms" in d:
                k = d["kernel_ms"]
                f.write(f"| {d['config']} | {d['label_maps']} | {d['mean_K']:.0f} | {1e3 * k['prep']:.0f} | {1e3 * k['narrow']:.0f} | {1e3 * k['big']:.0f} | "
                        f"{d['images_per_sec']:.3g} | {d['rewards_per_sec']:.3g} | {d['algorithmic_GBps']:.0f} | {d['max_rel_err_vs_oracle']:.1e} |\n")
        f.write("\nSlices of the same batches (`icr_slices_plan` + `icr_slices_emit`); HBM bytes = every input plane once + 4 B written per bbox pixel:\n\n"
                "| config | slices | packed bytes | plan us | emit us | HBM GB/s | fraction of the copy peak |\n|---|---|---|---|---|---|---|\n")
        for d in lines:
            if "emit_ms" in d:
                f.write(f"| {d['config']} | {d['segments']} | {d['slice_bytes'] / 1e6:.1f} MB | {1e3 * d['plan_ms']:.0f} | {1e3 * d['emit_ms']:.0f} | "
                        f"{d['emit_hbm_GBps']:.0f} | {d['emit_frac_of_hbm_peak']:.2f} |\n")
        f.write("\nDecoder (`icr_reassemble_map` one map at a time over 64 maps; `icr_reassemble_batch` all maps in one launch):\n\n"
                "| config | us per map | batch ms | batch maps/s | batch HBM GB/s | fraction of the copy peak |\n|---|---|---|---|---|---|\n")
        for d in lines:
            if "ms_per_map" in d:
                f.write(f"| {d['config']} | {1e3 * d['ms_per_map']:.1f} | {d['batch_ms']:.3f} | {d['batch_maps_per_sec']:.3g} | {d['batch_hbm_GBps']:.0f} | "
                        f"{d['batch_frac_of_hbm_peak']:.2f} |\n")
    print("wrote", f"{tag}_configs.md")


if __name__ == "__main__":
    tag, pre = sys.argv[1], sys.argv[2]
    bench_lines(tag, pre)
    launches(tag, pre)
    configs(tag, pre)
