"""Recipe for oracle/_ref/: the checker's copy of the UNMODIFIED reference, for the GPU box.

TEST / MEASUREMENT INFRASTRUCTURE ONLY.  /root/reference exists in the build container and nowhere else; the
reference is 2 934 lines of pure Python with nothing to compile, so "building" it is placing its files where the
snapshot that travels to the GPU box carries them: oracle/_ref/ is git-ignored (the reference's sources never enter
this repository's history) and NOT gpurun-ignored, exactly like the built libb2asr.so.  bench.py's `--impl reference`
arm and its `cpu_baseline` leg import it through oracle/reference_loader.py (by file path) and time the reference's
own Seq2Seq.forward + losses + backward + optimizer step (reference src/solver.py:132-182) on the host cores.
Nothing in the product package reads this directory.

    python -m oracle.make_ref        # or __graft_entry__.build()
"""
import filecmp
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
UPSTREAM = "/root/reference"
DEST = os.path.join(HERE, "_ref")
WHAT = ("src", "config", "main.py", "eval.py", "LICENSE")


def make():
    if not os.path.isdir(os.path.join(UPSTREAM, "src")):
        have = os.path.isfile(os.path.join(DEST, "src", "asr.py"))
        return "oracle/_ref: %s (no %s here)" % ("using the files already placed" if have else "absent", UPSTREAM)
    os.makedirs(DEST, exist_ok=True)
    n = 0
    for name in WHAT:
        src, dst = os.path.join(UPSTREAM, name), os.path.join(DEST, name)
        if os.path.isdir(src):
            for root, _, files in os.walk(src):
                rel = os.path.relpath(root, src)
                os.makedirs(os.path.join(dst, rel), exist_ok=True)
                for f in files:
                    if f.endswith((".py", ".yaml", ".md")):
                        a, b = os.path.join(root, f), os.path.join(dst, rel, f)
                        if not (os.path.exists(b) and filecmp.cmp(a, b, shallow=False)):
                            shutil.copyfile(a, b)
                        n += 1
        elif os.path.isfile(src):
            if not (os.path.exists(dst) and filecmp.cmp(src, dst, shallow=False)):
                shutil.copyfile(src, dst)
            n += 1
    return "oracle/_ref: %d files of the unmodified reference placed from %s" % (n, UPSTREAM)


if __name__ == "__main__":
    print(make())
