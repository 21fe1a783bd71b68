"""Reflection test runner with junit output (reference: py/kubeflow/tf_operator/test_runner.py
:20-196). It finds every ``test_*`` method of a suite class, runs each ``--num_trials`` times,
retries a failing attempt, and writes one junit XML per suite.

    python -m tf_operator_b200.testing.test_runner --suite all --artifacts_path /tmp/e2e

The cluster the reference reaches through a kubeconfig is started here, in process: API server,
operator and launcher on one box.
"""
from __future__ import annotations

import argparse
import inspect
import logging
import os
import sys
from typing import List, Optional, Type

from . import util

MAX_RETRIES = 3          # the reference retries 10x against a shared EKS cluster; one box is calmer


def add_common_args(parser: argparse.ArgumentParser) -> None:
    """test_runner.py:140-186."""
    parser.add_argument("--suite", default="all", help="suite name from testing.suites.SUITES, or all")
    parser.add_argument("--namespace", default="default")
    parser.add_argument("--tfjob_version", default="v1")
    parser.add_argument("--num_trials", type=int, default=1,
                        help="how many times each test method runs (same job name re-created)")
    parser.add_argument("--artifacts_path", default="", help="directory for junit_<suite>.xml")
    parser.add_argument("--junit_path", default="", help="explicit junit file (single suite)")
    parser.add_argument("--params", default="", help="comma separated k=v (kept for CLI parity)")
    parser.add_argument("--skip_tests", default="", help="comma separated test method names to skip")


def _test_methods(cls) -> List[str]:
    return sorted(n for n, f in inspect.getmembers(cls, predicate=inspect.isfunction)
                  if n.startswith("test_"))


def run_test(test_case, test_func, args) -> None:
    """One junit case = num_trials runs, each with up to MAX_RETRIES attempts
    (test_runner.py:45-98)."""
    def once():
        last: Optional[BaseException] = None
        for attempt in range(MAX_RETRIES):
            try:
                test_func()
                return
            except AssertionError:
                raise                       # a real verdict: do not retry it away
            except Exception as e:  # noqa: BLE001 - infrastructure flake: retry
                logging.exception("attempt %d of %s failed", attempt + 1, test_case.name)
                last = e
        raise last  # type: ignore[misc]

    def trials():
        for trial in range(max(1, args.num_trials)):
            logging.info("%s.%s trial %d", test_case.class_name, test_case.name, trial)
            once()
    try:
        util.wrap_test(trials, test_case)
    except Exception:  # noqa: BLE001 - recorded in the junit case
        logging.error("%s.%s FAILED: %s", test_case.class_name, test_case.name, test_case.failure)


def run_suite(suite_cls: Type, args, cluster) -> List[util.TestCase]:
    """Instantiate the suite and run each test method (test_runner.py:101-137)."""
    skip = {s for s in args.skip_tests.split(",") if s}
    cases = []
    for method in _test_methods(suite_cls):
        if method in skip:
            continue
        inst = suite_cls(args, cluster)
        tc = util.TestCase(class_name=suite_cls.__name__, name=method)
        run_test(tc, getattr(inst, method), args)
        cases.append(tc)
    return cases


def main(argv: Optional[List[str]] = None) -> int:
    from ..cluster import LocalCluster
    from ..launcher.kubelet import KubeletOptions
    from ..launcher.topology import synthetic
    from .fixtures import REPO
    from .suites import SUITES

    logging.basicConfig(level=logging.INFO, format="%(levelname)s|%(asctime)s|%(filename)s:%(lineno)d| %(message)s")
    parser = argparse.ArgumentParser(description="Run the TFJob e2e suites on the single-box cluster")
    add_common_args(parser)
    args = parser.parse_args(argv)
    names = list(SUITES) if args.suite == "all" else [s for s in args.suite.split(",") if s]
    unknown = [n for n in names if n not in SUITES]
    if unknown:
        parser.error(f"unknown suite(s) {unknown}; choose from {sorted(SUITES)}")
    failures = 0
    for name in names:
        cluster = LocalCluster(topology=synthetic(2), serve_http=True,
                               kubelet_options=KubeletOptions(cwd=REPO, restart_backoff=0.05))
        try:
            cases = run_suite(SUITES[name], args, cluster)
        finally:
            cluster.stop()
        failures += sum(1 for tc in cases if tc.failure)
        path = args.junit_path if (args.junit_path and len(names) == 1) else (
            os.path.join(args.artifacts_path, f"junit_{name}.xml") if args.artifacts_path else "")
        if path:
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
            util.create_junit_xml_file(cases, path)
        for tc in cases:
            print(f"{'FAIL' if tc.failure else 'ok  '} {tc.class_name}.{tc.name} ({tc.time:.2f}s)"
                  + (f"  {tc.failure}" if tc.failure else ""))
    return 1 if failures else 0


if __name__ == "__main__":
    sys.exit(main())
