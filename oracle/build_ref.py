"""TEST / BENCH INFRASTRUCTURE -- recipe for `oracle/_ref/`: the reference's own Python package staged for the
CPU baseline arm.

bennevans/mjrl is pure Python, so there is nothing to compile: this recipe stages the reference's
`mjrl/` package from where it lies (`/root/reference`, or $MJRL_REFERENCE) into `oracle/_ref/mjrl`, which is
git-ignored (it never enters the history) but not gpurun-ignored (it travels to the GPU box like a built .so),
so that `bench.py --impl reference` and the `cpu_baseline` leg time the REFERENCE's own
`process_samples.compute_returns / compute_advantages`, `NPG.train_from_paths` and `LinearBaseline.fit`
(`kind: "reference"`), not the port in oracle/mjrl_oracle.py.  Only bench.py and tests may use it; nothing
under mjrl_b200/ imports it (tests/test_abi_cpu.py).  `__graft_entry__.build()` runs this when the reference
checkout is present; without it (the GPU box) the already staged copy is used, and without that the port.

    python oracle/build_ref.py
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")


def build(reference=None, quiet=True):
    reference = reference or os.environ.get("MJRL_REFERENCE", "/root/reference")
    src = os.path.join(reference, "mjrl")
    if not os.path.isdir(src):
        return None
    dst = os.path.join(DEST, "mjrl")
    if os.path.isdir(dst):
        shutil.rmtree(dst)
    os.makedirs(DEST, exist_ok=True)
    # the Python modules of the update path; environments (MuJoCo XML, gym wrappers) are not needed
    shutil.copytree(src, dst, ignore=shutil.ignore_patterns("envs", "__pycache__", "*.pyc", "*.xml", "*.stl", "*.png"))
    with open(os.path.join(DEST, "SOURCE"), "w") as f:
        f.write("staged from %s by oracle/build_ref.py; not part of this repository's history\n" % src)
    if not quiet:
        print("staged", src, "->", dst)
    return dst


if __name__ == "__main__":
    out = build(quiet=False)
    sys.exit(0 if out else 1)
