"""Runs the reference's OWN unittest files (python/tests/test_*.py, read where they lie --
/root/reference in the build container, the staged oracle/_ref on the GPU box) against
THIS package: ``mpc_interface`` is aliased to ``mpc_interface_b200`` before they are
imported.  Prints "<run> <failures> <errors>".

    python tests/_run_reference_suite.py [--device] [<reference>/python/tests]

Without ``--device`` (no GPU: the build container) the two device entry points the host
classes call are stood in for by the test doubles the CPU suite uses everywhere (oracle
extend_matrices, NumpyBackend).  With ``--device`` nothing is replaced: every matrix the
reference's tests look at comes out of libmpci.so on the GPU (K1, the plan compiler's
DMMA products, the streaming kernels).
"""
import importlib
import importlib.util
import os
import sys
import types
import unittest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.dont_write_bytecode = True
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)


def main(test_dir, on_device=False):
    import mpc_interface_b200.device as device
    import mpc_interface_b200.tools as tools

    if on_device:
        device.require_gpu()
    else:
        from _numpy_backend import NumpyBackend
        from oracle import qp_oracle

        tools.extend_matrices = qp_oracle.extend_matrices
        device.DeviceBackend = NumpyBackend

    alias = types.ModuleType("mpc_interface")
    alias.__path__ = []
    sys.modules["mpc_interface"] = alias
    for sub in ("body", "combinations", "dynamics", "goal", "restrictions", "tools"):
        module = importlib.import_module("mpc_interface_b200." + sub)
        sys.modules["mpc_interface." + sub] = module
        setattr(alias, sub, module)
    # test_dynamics.py imports matplotlib for a plot helper it never calls
    for name in ("matplotlib", "matplotlib.pyplot", "matplotlib.colors"):
        if name not in sys.modules:
            try:
                importlib.import_module(name)
            except ImportError:
                sys.modules[name] = types.ModuleType(name)
                if "." in name:
                    setattr(sys.modules["matplotlib"], name.split(".")[1], sys.modules[name])

    test_dir = os.path.abspath(test_dir)
    os.chdir(test_dir)  # test_dynamics.py opens its pickle by a relative path
    suite = unittest.TestSuite()
    for f in sorted(os.listdir(test_dir)):
        if f.startswith("test_") and f.endswith(".py"):
            spec = importlib.util.spec_from_file_location("reference_" + f[:-3], os.path.join(test_dir, f))
            module = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(module)
            suite.addTests(unittest.defaultTestLoader.loadTestsFromModule(module))
    result = unittest.TextTestRunner(stream=sys.stderr, verbosity=2).run(suite)
    example = run_example(os.path.join(os.path.dirname(test_dir), "use_examples", "simple_functional_example"))
    print("example", "ok" if example else "FAILED")
    if on_device:  # proof that the kernels ran: the library is loaded and saw a device
        print("backend", "libmpci" if device._lib is not None and device.lib().mpci_device_count() > 0
              else "MISSING")
    print(result.testsRun, len(result.failures), len(result.errors))
    return 0 if result.wasSuccessful() and example else 1


def run_example(example_dir):
    """The reference's example formulation (use_examples/simple_functional_example/
    biped_formulation.py, unmodified) built on this package's classes: its QP of the
    default configuration and of the receding-horizon ticks 0, 11, 12, 13 against the
    golden vectors the reference itself produced (tests/golden/biped_c1, biped_ticks)."""
    import numpy as np

    sys.path.insert(0, example_dir)
    import biped_configuration
    from biped_formulation import formulate_biped

    def close(got, want):
        worst = 0.0
        for x, r in zip(got, want):
            if np.shape(x) != np.shape(r):
                return False
            worst = max(worst, float(np.max(np.abs(x - r)) / max(np.max(np.abs(r)), 1e-300)))
        return worst <= 1e-10

    golden = os.path.join(HERE, "golden")
    form = formulate_biped(biped_configuration)
    if not type(form).__module__.startswith("mpc_interface_b200"):
        return False
    g = np.load(os.path.join(golden, "biped_c1.npz"))
    ok = close(form.generate_all_qp_matrices(g["given"]), [g[k] for k in "AhQq"])
    t = np.load(os.path.join(golden, "biped_ticks.npz"))
    n = biped_configuration.step_samples
    step_times = np.array([(i + 1) * n - 1 for i in range(form.domain["Ds_x"])])
    step_count = 0
    for tick in range(14):
        form.update(step_times=step_times, step_count=step_count)
        if tick in (0, 11, 12, 13):
            got = form.generate_all_qp_matrices(t["t%d_given" % tick])
            ok = ok and close(got, [t["t%d_%s" % (tick, k)] for k in "AhQq"])
        step_times = step_times - 1
        if step_times[0] == -1:
            step_times = step_times + n
            step_count += 1
    return ok


if __name__ == "__main__":
    argv = [a for a in sys.argv[1:] if a != "--device"]
    sys.exit(main(argv[0] if argv else "/root/reference/python/tests", "--device" in sys.argv[1:]))
