#!/usr/bin/env python
"""Side-by-side table of serving sweeps (the JSON files written by scripts/benchmark_serving.py and
baseline/standin_serve.py): output tokens/s, p50 / p99 TTFT, mean inter-token latency and p50 end-to-end latency per
concurrency level, plus the ratio of each file to the first one. The serving counterpart of
scripts/compare_training.py (the reference declares load testing, README.md:17, but has no report tool).

    python scripts/compare_serving.py profiles/vllm_standin_tp1_r1_unique_prompts.json profiles/serving_bench_tp1_r1_v6_async.json
"""
import argparse
import json
import os


def load(path: str):
    d = json.load(open(path))
    name = d.get("impl") or f"dlti_b200 tp={d.get('tp')}"
    if "frontend" in d:
        name += f" ({'async' if d.get('async_scheduling') else 'sync'}, {d['frontend']} front-end)"
    return name, {int(l["concurrency"]): l for l in d["levels"]}


def table(paths: list[str]) -> str:
    runs = [load(p) for p in paths]
    levels = sorted(set().union(*[set(r[1]) for r in runs]))
    head = "| users | " + " | ".join(f"{n}<br>tok/s · TTFT p50/p99 ms · ITL ms · e2e p50 ms" for n, _ in runs)
    if len(runs) > 1:
        head += " | tok/s vs first"
    lines = [head + " |", "|" + "---|" * (len(runs) + 1 + (len(runs) > 1))]
    for c in levels:
        cells, ratios = [], []
        base = runs[0][1].get(c)
        for _, lv in runs:
            l = lv.get(c)
            if l is None:
                cells.append("-")
                continue
            cells.append(f"{l['output_tokens_per_s']:,.0f} · {l['ttft_p50_ms']:.0f}/{l['ttft_p99_ms']:.0f} · "
                         f"{l['itl_mean_ms']:.2f} · {l['e2e_p50_ms']:.0f}")
            if base is not None and l is not base:
                ratios.append(f"{l['output_tokens_per_s'] / base['output_tokens_per_s']:.2f}x")
        row = f"| {c} | " + " | ".join(cells)
        if len(runs) > 1:
            row += " | " + ", ".join(ratios)
        lines.append(row + " |")
    return "\n".join(lines)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("files", nargs="+")
    ap.add_argument("--out", default=None, help="also write the markdown table here")
    a = ap.parse_args()
    md = table(a.files)
    md += "\n\nfiles: " + ", ".join(os.path.basename(f) for f in a.files) + "\n"
    print(md)
    if a.out:
        open(a.out, "w").write(md)


if __name__ == "__main__":
    main()
