#!/usr/bin/env python3
"""Regenerate docs/parameters.md from the option tables in hetseq_9cme_b200/options.py."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hetseq_9cme_b200 import options  # noqa: E402

SECTIONS = [("General", "_GENERAL"), ("Dataset / data loading", "_DATA_COMMON"), ("Training-only data flags", "_DATA_TRAIN"),
            ("BERT files", "_BERT_FILES"), ("Fine-tuning (NER / EL) files", "_FINE_TUNE"), ("Task-specific", "_TASK_ROWS"),
            ("Optimisation", "_OPT_COMMON"), ("Per optimizer", "_OPT_ROWS"), ("Per learning-rate scheduler", "_SCHED_ROWS"),
            ("Checkpointing", "_CHECKPOINT"), ("B200 engine (new)", "_ENGINE")]


def rows_to_md(rows):
    out = ["| flag | default | meaning |", "|---|---|---|"]
    for names, kw in rows:
        default = "off" if kw.get("action") == "store_true" else kw.get("default", "")
        text = (kw.get("help") or "").replace("|", "/").replace("%%", "%")
        if kw.get("choices"):
            text += " (choices: " + ", ".join(str(c) for c in kw["choices"]) + ")"
        out.append("| `%s` | `%s` | %s |" % (", ".join(names), default, text))
    return "\n".join(out)


def table(name):
    obj = getattr(options, name)
    if isinstance(obj, dict):
        return "\n".join("**%s**\n\n%s\n" % (k, rows_to_md(v)) for k, v in obj.items())
    return rows_to_md(obj)


def main():
    parts = ["# Command-line parameters\n",
             "Generated from `hetseq_9cme_b200/options.py` by `python tools/gen_param_docs.py`.  The flag set mirrors the\n"
             "reference's `hetseq/options.py` group by group (general, dataset, distributed, optimisation, checkpoint); rows\n"
             "that depend on `--task / --optimizer / --lr-scheduler` are added once those are known; one group is new.\n"]
    for title, name in SECTIONS:
        parts.append("## %s\n\n%s\n" % (title, table(name)))
    parts.append("Distributed flags (`--distributed-world-size`, `--distributed-gpus`, `--distributed-rank`, `--distributed-backend`,\n"
                 "`--distributed-init-method`, `--device-id`, `--bucket-cap-mb`, `--fast-stat-sync`, `--find-unused-parameters`)\n"
                 "are described with the launch modes in `docs/distribute.md`.\n")
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "docs", "parameters.md")
    with open(path, "w") as f:
        f.write("\n".join(parts))
    print("wrote", path)


if __name__ == "__main__":
    main()
