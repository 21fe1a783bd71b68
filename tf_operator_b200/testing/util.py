"""Shared utilities of the e2e harness (reference: py/kubeflow/tf_operator/util.py and
test_util.py): subprocess helper, service-proxy requests to a replica's fake workload, timing
and junit bookkeeping.
"""
from __future__ import annotations

import logging
import subprocess
import time
import urllib.error
import urllib.request
from typing import Dict, List, Optional
from xml.etree import ElementTree as ET


class TimeoutError_(TimeoutError):
    """util.py:68 - raised by the wait helpers."""


def run(command: List[str], cwd: Optional[str] = None, env: Optional[Dict[str, str]] = None,
        dryrun: bool = False) -> str:
    """Run a subprocess, log and return its combined output (util.py:28-65)."""
    logging.info("Running: %s \ncwd=%s", " ".join(command), cwd)
    if dryrun:
        return ""
    out = subprocess.run(command, cwd=cwd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                         text=True)
    for line in out.stdout.splitlines():
        logging.info(line)
    if out.returncode != 0:
        raise subprocess.CalledProcessError(out.returncode, command, out.stdout)
    return out.stdout


def send_request(api_url: str, namespace: str, master_host: str, rpc: str, params: Dict[str, object],
                 port: int = 2222, retries: int = 200, interval: float = 0.05) -> str:
    """GET <apiserver>/api/v1/namespaces/<ns>/services/<svc>:<port>/proxy/<rpc>?<params>
    - the route the reference uses to reach a replica's test-server (util.py:108-139; its EKS
    bearer token has no counterpart on one box). Retries while the replica is still binding."""
    query = "&".join(f"{k}={v}" for k, v in params.items())
    url = f"{api_url}/api/v1/namespaces/{namespace}/services/{master_host}:{port}/proxy/{rpc}"
    if query:
        url += "?" + query
    last: Optional[Exception] = None
    for _ in range(retries):
        try:
            with urllib.request.urlopen(url, timeout=5) as r:
                return r.read().decode()
        except (urllib.error.URLError, ConnectionError, OSError) as e:
            last = e
            time.sleep(interval)
    raise RuntimeError(f"request {url} failed: {last}")


# ---- junit (test_util.py:16-117) --------------------------------------------------------------
class TestCase:
    def __init__(self, class_name: str = "", name: str = ""):
        self.class_name = class_name
        self.name = name
        self.time: Optional[float] = None
        self.failure: Optional[str] = None

    __test__ = False     # not a pytest class


class TestSuite:
    def __init__(self, class_name: str):
        self._cases: Dict[str, TestCase] = {}
        self._class_name = class_name

    def create(self, name: str) -> TestCase:
        if name in self._cases:
            raise ValueError(f"TestSuite already has a test named {name}")
        self._cases[name] = TestCase(self._class_name, name)
        return self._cases[name]

    def get(self, name: str) -> TestCase:
        if name not in self._cases:
            raise ValueError(f"No TestCase named {name}")
        return self._cases[name]

    def __iter__(self):
        return iter(self._cases.values())

    __test__ = False


def wrap_test(test_func, test_case: TestCase) -> None:
    """Run test_func, record its time and failure in test_case, re-raise (test_util.py:119-151)."""
    start = time.time()
    try:
        test_func()
    except subprocess.CalledProcessError as e:
        test_case.failure = f"Subprocess failed;\n{e.output}"
        raise
    except Exception as e:  # noqa: BLE001
        test_case.failure = f"Test failed; {type(e).__name__}: {e}"
        raise
    finally:
        test_case.time = time.time() - start


def create_xml(test_cases) -> ET.ElementTree:
    """junit document for a list of TestCase (test_util.py:64-117)."""
    cases = list(test_cases)
    total_time = sum(tc.time or 0 for tc in cases)
    failures = sum(1 for tc in cases if tc.failure)
    root = ET.Element("testsuite", {"failures": str(failures), "tests": str(len(cases)),
                                    "time": f"{total_time}"})
    for tc in cases:
        attrib = {"classname": tc.class_name, "name": tc.name}
        if tc.time is not None:
            attrib["time"] = f"{tc.time}"
        # a case without a time never ran: junit has no "not run", so it counts as failed
        e = ET.SubElement(root, "testcase", attrib)
        if tc.failure or tc.time is None:
            ET.SubElement(e, "failure").text = tc.failure or "Test was not run"
    return ET.ElementTree(root)


def create_junit_xml_file(test_cases, output_path: str) -> None:
    create_xml(test_cases).write(output_path, encoding="utf-8", xml_declaration=True)
    logging.info("wrote junit file %s", output_path)


def get_num_failures(xml_string: str) -> int:
    """test_util.py:154-167."""
    return int(ET.fromstring(xml_string).attrib.get("failures", 0))
