"""Builds tests/cpp/test_gecko.cpp (the reference's C++ test cases for
gecko::tools::extend_matrices) against include/mpci_gecko.hh + libmpci.so."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "mpc_interface_b200", "csrc")
OUT = os.path.join(ROOT, "tests", "cpp", "build", "test_gecko")


def build():
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    subprocess.run(
        ["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
         os.path.join(ROOT, "tests", "cpp", "test_gecko.cpp"), "-o", OUT,
         "-L", CSRC, "-lmpci", "-Wl,-rpath," + CSRC],
        check=True)
    return OUT


def run(*args):
    return subprocess.run([build(), *args], capture_output=True, text=True, timeout=300)
