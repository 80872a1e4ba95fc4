#!/usr/bin/env python3
"""Stages the UNMODIFIED reference (pure Python) into oracle/_ref/ -- TEST INFRASTRUCTURE.

    python oracle/make_ref.py [/root/reference]

The reference's path is Python (python/mpc_interface/*.py, 1.9 kLoC, numpy + scipy +
sympy); it has no build step and is not pip-installable offline (its install is a CMake
rule, python/CMakeLists.txt:10-12), so "building" oracle/_ref is a byte-for-byte copy of
`python/` -- the package, the example formulation and the reference's own unit tests with
their pickle fixture -- plus MANIFEST.json (sha256 of every file, so a test can prove the
staged tree is the unmodified reference).  oracle/_ref/ is git-ignored (no reference
source enters the history) but not gpurun-ignored: it travels to the GPU box, where it is
(1) the timed CPU arm of bench.py (`cpu_baseline.kind = "reference"`) and (2) a live
oracle for the `-m gpu` parity tests.  Called by __graft_entry__.build() whenever
/root/reference is present; on the GPU box the staged copy is used as it arrived.
"""
import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")


def sha256(path):
    with open(path, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def stage(reference="/root/reference"):
    """Copies <reference>/python to oracle/_ref/python; returns the manifest or None
    when the reference is not present (the GPU box)."""
    src = os.path.join(reference, "python")
    if not os.path.isdir(os.path.join(src, "mpc_interface")):
        return None
    dst = os.path.join(DEST, "python")
    if os.path.isdir(DEST):
        shutil.rmtree(DEST)
    # contents only: the checkout is read-only, the staged copy must stay removable
    for base, dirs, files in os.walk(src):
        dirs[:] = [d for d in dirs if d != "__pycache__"]
        out = os.path.join(dst, os.path.relpath(base, src))
        os.makedirs(out, exist_ok=True)
        for name in files:
            if not name.endswith(".pyc"):
                shutil.copyfile(os.path.join(base, name), os.path.join(out, name))
    manifest = {"source": src, "files": {}}
    for base, _, files in sorted(os.walk(dst)):
        for name in sorted(files):
            path = os.path.join(base, name)
            rel = os.path.relpath(path, dst)
            manifest["files"][rel] = sha256(path)
            assert manifest["files"][rel] == sha256(os.path.join(src, rel)), rel
    with open(os.path.join(DEST, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    # hashes only (no reference source) are tracked, so that a staged tree can be proven
    # to be the unmodified reference wherever it is used
    with open(os.path.join(HERE, "REF_MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1, sort_keys=True)
    return manifest


def verify(path=None):
    """True when every file of the staged tree has the tracked sha256 (and none is missing)."""
    path = os.path.join(DEST, "python") if path is None else path
    with open(os.path.join(HERE, "REF_MANIFEST.json")) as f:
        want = json.load(f)["files"]
    return all(os.path.isfile(os.path.join(path, rel)) and sha256(os.path.join(path, rel)) == digest
               for rel, digest in want.items())


if __name__ == "__main__":
    done = stage(sys.argv[1] if len(sys.argv) > 1 else "/root/reference")
    if done is None:
        print("reference not found; oracle/_ref left as it is")
        sys.exit(0 if os.path.isdir(DEST) else 1)
    print("staged {} files into {}".format(len(done["files"]), DEST))
