"""E2E harness: the reference's `py/kubeflow/tf_operator`
test helpers, reflection runner with junit XML, job fixtures and suites, re-expressed against
the single-box cluster (in-memory API server + operator + process launcher) instead of an
Argo workflow on EKS.

    python -m tf_operator_b200.testing.test_runner --suite simple_tfjob_tests --junit_path out.xml
"""
