#!/usr/bin/env python3
"""Stage an UNMODIFIED copy of the reference's Python module under oracle/_ref/ (git-ignored).

Test infrastructure only.  The reference (andrewginns/quorum) is pure Python, so "building" it is a
file copy: src/quorum/{__init__,oai_proxy}.py are copied byte for byte from /root/reference, and a
small concatenate-only config.yaml (written here, not copied) is put where the module looks for it
(`Path(__file__).parent.parent.parent / "config.yaml"`, reference src/quorum/oai_proxy.py:46).
Importing the module writes logs/aggregation.log next to that config (oai_proxy.py:20-37), which is
why it is never imported from /root/reference itself.

Nothing under oracle/_ref/ is committed; it travels to the GPU box with the gpurun snapshot so that
`bench.py --impl reference` can time the real reference there.  When /root/reference is absent
(GPU box) this script is a no-op and whatever is already staged is used.
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/src/quorum"
DST = os.path.join(HERE, "_ref")

CONFIG = """\
primary_backends:
  - name: "B1"
    url: "http://test1.example.com/v1"
    model: "m1"
  - name: "B2"
    url: "http://test2.example.com/v1"
    model: "m2"
iterations:
  aggregation:
    strategy: concatenate
strategy:
  concatenate:
    separator: "\\n-------------\\n"
    hide_intermediate_think: true
    hide_final_think: true
    thinking_tags: ["think", "reason", "reasoning", "thought"]
    skip_final_aggregation: false
settings:
  timeout: 60
"""


def stage() -> bool:
    if not os.path.isdir(SRC):
        return os.path.isfile(os.path.join(DST, "src", "quorum", "oai_proxy.py"))
    pkg = os.path.join(DST, "src", "quorum")
    os.makedirs(pkg, exist_ok=True)
    for name in ("__init__.py", "oai_proxy.py"):
        shutil.copyfile(os.path.join(SRC, name), os.path.join(pkg, name))
    with open(os.path.join(DST, "config.yaml"), "w") as f:
        f.write(CONFIG)
    return True


if __name__ == "__main__":
    ok = stage()
    print("oracle/_ref staged" if ok else "reference not available", file=sys.stderr)
    sys.exit(0)
