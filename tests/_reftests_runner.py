"""Subprocess body of tests/test_reference_suite_gpu.py: runs the REFERENCE's own unit tests for the hot path
(/root/reference/tests/L0/run_amp and run_optimizers, staged unmodified, as a bytecode archive, in oracle/_ref/pyref_tests.zip by
oracle/build_ref.sh) with the reference's own Python package (oracle/_ref/pyref.zip) over one of two back ends:

  python tests/_reftests_runner.py ours <out.json> <dir/test_file.py> ...   amp_C = this repo's shim over libb200mt.so
  python tests/_reftests_runner.py ref  <out.json> <dir/test_file.py> ...   amp_C = the reference's kernels (amp_C_ref.so)

Writes {test id: "ok" | "fail: ..." | "error: ..." | "skip: ..."} as JSON.
"""
import importlib
import json
import os
import sys
import unittest

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from _refpy_worker import ROOT, setup   # noqa: E402

TESTS = os.path.join(ROOT, "oracle", "_ref", "pyref_tests.zip")


class Result(unittest.TextTestResult):
    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self.outcomes = {}

    def startTest(self, test):
        # Several of the reference's tests draw unseeded random data and models (test_checkpointing.test_loss_scale_decrease
        # needs its random model to overflow twice in a row): both back ends must see the SAME draws, test by test, or the
        # differential contract compares two different experiments.  2809 is the seed the reference's own tests use where they seed.
        import torch
        torch.manual_seed(2809)
        torch.cuda.manual_seed_all(2809)
        super().startTest(test)

    def addSuccess(self, test):
        super().addSuccess(test); self.outcomes[test.id()] = "ok"

    def addFailure(self, test, err):
        super().addFailure(test, err); self.outcomes[test.id()] = "fail: " + self._exc_info_to_string(err, test)[-1500:]

    def addError(self, test, err):
        super().addError(test, err); self.outcomes[test.id()] = "error: " + self._exc_info_to_string(err, test)[-1500:]

    def addSkip(self, test, reason):
        super().addSkip(test, reason); self.outcomes[test.id()] = "skip: " + reason


def main():
    backend, out, files = sys.argv[1], sys.argv[2], sys.argv[3:]
    setup(backend)
    outcomes = {}
    for rel in files:
        d, f = os.path.split(rel)
        sys.path.insert(0, os.path.join(TESTS, d))          # the tests do `from utils import common_init`
        # the staged tests are a bytecode archive (oracle/build_ref.sh): import the module through zipimport
        name = os.path.splitext(f)[0]
        sys.modules.pop("utils", None)                        # each test directory has its own utils module
        mod = importlib.import_module(name)
        suite = unittest.defaultTestLoader.loadTestsFromModule(mod)
        runner = unittest.TextTestRunner(resultclass=Result, verbosity=1, stream=sys.stderr)
        r = runner.run(suite)
        outcomes.update(r.outcomes)
        sys.path.pop(0)
    with open(out, "w") as fh:
        json.dump(outcomes, fh, indent=1)
    print("reference tests over %s: %d ok of %d" % (backend, sum(v == "ok" for v in outcomes.values()), len(outcomes)))


if __name__ == "__main__":
    main()
