"""Version info (pkg/version/version.go:21-43)."""
import platform
import subprocess
import sys

VERSION = "v0.1.0-alpha"
GIT_SHA = "Not provided."


def _git_sha() -> str:
    try:
        import os
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        return subprocess.run(["git", "-C", root, "rev-parse", "--short", "HEAD"],
                              capture_output=True, text=True, timeout=2).stdout.strip() or GIT_SHA
    except Exception:  # noqa: BLE001
        return GIT_SHA


def info() -> dict:
    return {"version": VERSION, "git_sha": _git_sha(), "python": sys.version.split()[0],
            "platform": f"{platform.system().lower()}/{platform.machine()}"}


def print_version_and_exit(api_version: str = "v1") -> None:
    i = info()
    print(f"Version: {i['version']}")
    print(f"Git SHA: {i['git_sha']}")
    print(f"Python Version: {i['python']}")
    print(f"OS/Arch: {i['platform']}")
    print(f"API Version: {api_version}")
    sys.exit(0)
