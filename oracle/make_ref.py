#!/usr/bin/env python
"""Recipe for oracle/_ref/: the UNMODIFIED reference watcher, staged so that it can run where /root/reference is absent.

TEST INFRASTRUCTURE.  The reference (highreso-gpu/k8s-watcher) is a pure-Python script tree with no build system
and no C sources, so "building" it means staging its few source files: watcher/*.py and config/*.yaml are copied
byte for byte from /root/reference into oracle/_ref/ (git-ignored, so the reference's sources never enter this
repository's history; NOT gpurun-ignored, so the staged copy travels to the GPU box).  Only bench.py's
`--impl reference` / `cpu_baseline` legs and the tests may import what lands there -- as the thing that is timed
beside the GPU path or checked against, never as part of the product.

    python oracle/make_ref.py           # no-op (exit 0) when /root/reference is not there
"""
import filecmp
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
DST = os.path.join(HERE, "_ref")
FILES = ["watcher/__init__.py", "watcher/pod_watcher.py", "watcher/clusterapi_client.py",
         "config/base.yaml", "config/development.yaml", "config/production.yaml", "config/staging.yaml"]


def main() -> int:
    if not os.path.isdir(REF):
        print("oracle/_ref: %s not present, keeping what is staged (%s)" % (REF, "present" if os.path.isdir(DST) else "nothing"))
        return 0
    for rel in FILES:
        src, dst = os.path.join(REF, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if not (os.path.exists(dst) and filecmp.cmp(src, dst, shallow=False)):
            shutil.copyfile(src, dst)
    print("oracle/_ref: staged %d files of the unmodified reference" % len(FILES))
    return 0


def load_reference():
    """-> the reference's watcher.pod_watcher module from oracle/_ref (under the stand-in kubernetes package of
    tests/fake_kubernetes.py), or None when nothing is staged.  Its config paths are CWD-relative
    (pod_watcher.py:22,25): construct its PodWatcher with oracle/_ref as the working directory."""
    if not os.path.exists(os.path.join(DST, "watcher", "pod_watcher.py")):
        return None
    sys.path.insert(0, os.path.join(os.path.dirname(HERE), "tests"))
    import fake_kubernetes as fk
    fk.install()
    if DST not in sys.path:
        sys.path.insert(0, DST)
    sys.dont_write_bytecode = True
    import importlib
    return importlib.import_module("watcher.pod_watcher")


if __name__ == "__main__":
    sys.exit(main())
