"""Regenerate the measured-numbers tables of README.md and BASELINE.md from the committed bench lines
(profiles/bench_r2_final.json = the default `python bench.py` line, profiles/bench_r2_final_n{2,4,8}.json = the
torchrun lines, profiles/bench_r2_final_reference.json = `--impl reference`).  Text between the markers
`<!-- measured:begin -->` / `<!-- measured:end -->` is replaced."""
import json
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
P = ROOT / "profiles"


def load(name):
    f = P / name
    return json.loads(f.read_text()) if f.exists() and f.read_text().strip() else None


def pct(x):
    return "—" if x is None else f"{100 * x:.1f} %"


def table() -> str:
    d = load("bench_r2_final.json")
    if d is None:
        return "(no bench line committed yet)\n"
    out = []
    r, ra, rs = d.get("roofline_score_select_pack", d["roofline"]), d["roofline_attention"], d["roofline_step"]
    one = d["roofline"] if "roofline_score_select_pack" in d else d.get("roofline_one_launch_layer")
    out.append(f"Measured on 1 x B200 (`python bench.py`, CUDA-graph replay, clocks {d['clocks']['sm_mhz']:.0f} MHz, "
               f"no throttle reasons), roofline against the measured copy peak {r['peak']:.0f} GB/s "
               f"(`MEASURED_PEAKS.json`) and, in brackets, against the nominal 8 TB/s:\n")
    out.append("| workload | tok/s (`value`) | ms/step | tok/s end to end (`e2e`) | score+select+pack launch | attention launch | whole step |")
    out.append("|---|---|---|---|---|---|---|")
    one_txt = (f"; one-launch layer {one['avg_launch_ms'] * 1e3:.1f} us, {pct(one['frac'])} ({pct(one.get('frac_of_nominal_8tbs'))})"
               if one and "roofline_score_select_pack" in d else "")
    out.append(f"| cfg5 Qwen3-32B shape, bs16, ctx 128K, budget 4096, skip 0 | **{d['value']:.1f}** | {d['ms_per_step']:.3f} | "
               f"**{d['e2e']['value']:.1f}** | {r['avg_launch_ms'] * 1e3:.1f} us, {pct(r['frac'])} ({pct(r['frac_of_nominal_8tbs'])}){one_txt} | "
               f"{ra['avg_launch_ms'] * 1e3:.1f} us, {pct(ra['frac'])} ({pct(ra['frac_of_nominal_8tbs'])}) | "
               f"{pct(rs['frac'])} ({pct(rs['frac_of_nominal_8tbs'])}) |")
    names = {"cfg5_skip2": "cfg5, quest_skip_layers = 2 (two dense 128K layers)", "cfg3": "cfg3 Qwen2.5-7B shape, bs32, ctx 128K, budget 4096",
             "cfg2": "cfg2 Llama-3.1-8B shape, bs1, ctx 32K, budget 2048", "cfg4": "cfg4 Llama-3.1-8B shape, H2O, bs64, ctx 32K, budget 2048"}
    for key, name in names.items():
        o = d.get("other_workloads", {}).get(key)
        if not o or "value" not in o:
            continue
        att = (f"{o['attention_launch_ms'] * 1e3:.1f} us, {pct(o.get('roofline_attention_frac'))}"
               if o.get("attention_launch_ms") else "—")
        dom = f"{o['roofline_launch_ms'] * 1e3:.1f} us, {pct(o.get('roofline_frac'))}"
        if "score_select_pack_launch_ms" in o:  # one-launch layers: the halves timed alone, then the layer kernel
            dom = (f"{o['score_select_pack_launch_ms'] * 1e3:.1f} us, {pct(o['score_select_pack_frac'])}; one-launch layer "
                   f"{o['roofline_launch_ms'] * 1e3:.1f} us, {pct(o.get('roofline_frac'))}")
        if key == "cfg4":
            dom, att = "—", dom + " (scored attention)"
        out.append(f"| {name} | {o['value']:.1f} | {o['ms_per_step']:.3f} | {o['e2e']:.1f} | {dom} | {att} | {pct(o.get('roofline_step_frac'))} |")
    if one:
        used = "the path the timed step uses" if "roofline_score_select_pack" in d else "not used at this shape: the two launches are faster"
        out.append(f"\nOne-launch layer at cfg5 (store + score + select + pack + attention, `svb200_quest_decode_layer`; {used}): "
                   f"{one['avg_launch_ms'] * 1e3:.1f} us per layer, {pct(one['frac'])} ({pct(one.get('frac_of_nominal_8tbs'))}) "
                   f"for the {one['bytes_per_launch'] / 1e6:.0f} MB of both halves; the two launches, each alone: "
                   f"{one['two_launches_ms'] * 1e3:.1f} us.")
    o = d.get("other_workloads", {}).get("cfg5_skip2") or {}
    if o.get("dense_layer_launch_ms"):
        out.append(f"\nDense skip layer at cfg5 (8.6 GB per launch): {o['dense_layer_launch_ms']:.3f} ms, {pct(o.get('dense_layer_frac'))} of the measured copy peak.")
    inc, cpu = d.get("incumbent") or {}, d.get("cpu_baseline") or {}
    if "value" in inc:
        out.append(f"\nThe reference's own GPU path on the same B200 (`--impl incumbent`: torch selection + Triton stage 1/2, unmodified, from "
                   f"`oracle/_ref`): {inc['value']:.1f} tok/s at cfg5 ({inc['select_ms_per_layer']:.2f} ms selection + "
                   f"{inc['attention_ms_per_layer']:.3f} ms attention per layer) -> **{d['value'] / inc['value']:.1f}x**.")
    if "value" in cpu:
        c1 = cpu.get("cfg1_native") or {}
        out.append(f"The reference's CPU path on the box's {cpu['cores']} host cores (`cpu_baseline`, kind `{cpu['kind']}`): "
                   f"{cpu['value']:.3f} tok/s at cfg5" + (f"; cfg1 natively {c1['value']:.1f} tok/s" if "value" in c1 else "") + ".")
    ref = load("bench_r2_final_reference.json")
    if ref and "value" in ref:
        out.append(f"`bench.py --impl reference`: {ref['value']:.3f} tok/s ({ref['ms_per_step'] / 1e3:.1f} s per cfg5 step).")
    rows = []
    for n in (2, 4, 8):
        x = load(f"bench_r2_final_n{n}.json")
        if x:
            c = x.get("collective") or {}
            rows.append(f"| {n} | {x['value']:.1f} | {x['ms_per_step']:.3f} | {x['e2e']['value']:.1f} | {x['value'] / d['value'] / n:.2f} | "
                        f"{x.get('roofline_score_select_pack', x['roofline'])['avg_launch_ms'] * 1e3:.1f} | {x['roofline_attention']['avg_launch_ms'] * 1e3:.1f} | "
                        f"{c.get('us_per_layer', float('nan')):.1f} | {c.get('other_us_per_layer', float('nan')):.1f} |")
    if rows:
        out.append("\nTensor parallel (cfg5, KV heads sharded, one process per GPU, NVLink push epilogue, max over ranks):\n")
        out.append("| GPUs | tok/s | ms/step | e2e tok/s | efficiency vs N=1 | score+select us | attention us | collective us/layer | gaps + slowest rank us/layer |")
        out.append("|---|---|---|---|---|---|---|---|---|")
        out.append(f"| 1 | {d['value']:.1f} | {d['ms_per_step']:.3f} | {d['e2e']['value']:.1f} | 1.00 | {r['avg_launch_ms'] * 1e3:.1f} | {ra['avg_launch_ms'] * 1e3:.1f} | — | — |")
        out.extend(rows)
    return "\n".join(out) + "\n"


def main():
    block = "<!-- measured:begin -->\n" + table() + "<!-- measured:end -->"
    for name in ("README.md", "BASELINE.md"):
        f = ROOT / name
        s = f.read_text()
        if "<!-- measured:begin -->" in s:
            s = re.sub(r"<!-- measured:begin -->.*?<!-- measured:end -->", lambda m: block, s, flags=re.S)
        elif "MEASURED_TABLE" in s:
            s = s.replace("MEASURED_TABLE", block)
        else:
            s = s.rstrip("\n") + "\n\n## 6. Measured on this pool's B200 (round 2)\n\n" + block + "\n"
        f.write_text(s)
        print("updated", name)


if __name__ == "__main__":
    main()
