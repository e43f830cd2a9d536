#!/usr/bin/env python
"""Copies what a round-2 GPU visit left in gpurun_out/ into tracked files under profiles/ (bench lines as one JSON object per
file, test-log tails, the attention micro-benchmark) and prints a one-screen summary.

    python scripts/collect_r2.py r2a        # prefix of the visit's files in gpurun_out/ and of the files written
"""
import glob
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GO, PROF = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")


def last_json_line(path):
    try:
        lines = [l for l in open(path, errors="replace") if l.lstrip().startswith("{")]
        return json.loads(lines[-1]) if lines else None
    except Exception:
        return None


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r2a"
    os.makedirs(PROF, exist_ok=True)
    summary = {}
    for path in sorted(glob.glob(os.path.join(GO, f"{tag}_bench*"))):
        if path.endswith(".err"):
            continue
        d = last_json_line(path)
        name = os.path.splitext(os.path.basename(path))[0]
        if d is None:
            print(name, ": no JSON line")
            continue
        with open(os.path.join(PROF, name + ".json"), "w") as f:
            f.write(json.dumps(d) + "\n")
        if "roofline" in d:
            st = d.get("stage_ms_per_step", {})
            summary[name] = {"ms_per_step": round(d["ms_per_step"], 3), "value": round(d["value"]), "e2e": round(d["e2e"]["value"]),
                             "e2e_sync": round(d["e2e"].get("synchronous_call_value", 0)), "gemm_frac": round(d["roofline"]["frac"], 3),
                             "attn_frac": round(d["roofline_attention"]["frac"], 3), "attn_ms": round(d["roofline_attention"]["ms_per_step"], 3),
                             "mel_frac": round(d["roofline_mel"]["frac"], 3), "stem_ms": round(st.get("audio_encoding.conv_chunk", 0), 3),
                             "xfmr_ms": round(st.get("audio_encoding.transformer", 0), 3), "launches": d.get("gpu_launches"),
                             "value_checked": d.get("value_checked"), "corpus_audio_s_per_s": (d.get("corpus") or {}).get("audio_s_per_s"),
                             "clocks": d.get("clocks")}
        else:
            summary[name] = {"value": d.get("value"), "impl": d.get("impl")}
    tails = {}
    for path in sorted(glob.glob(os.path.join(GO, f"{tag}_*pytest*.log"))) + sorted(glob.glob(os.path.join(GO, f"{tag}_attn_bench.log"))):
        lines = open(path, errors="replace").read().strip().splitlines()
        tails[os.path.basename(path)] = lines[-14:] if "attn_bench" in path else lines[-4:]
    with open(os.path.join(PROF, f"{tag}_summary.json"), "w") as f:
        json.dump({"bench": summary, "logs": tails}, f, indent=1)
    for k, v in summary.items():
        print(k, json.dumps(v))
    for k, v in tails.items():
        print("==", k)
        print("\n".join(v))


if __name__ == "__main__":
    main()
