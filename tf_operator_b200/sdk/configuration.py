"""``Configuration``: where the API server is and how to talk to it (reference: the swagger
runtime's sdk/python/kubeflow/tfjob/configuration.py - host, api_key, verify_ssl, logging knobs).
Generic runtime piece; the job clients do not need it (they take a transport), it is here so
code written against the reference package's ``ApiClient(Configuration())`` keeps working.
"""
from __future__ import annotations

import copy
import logging
import os
import sys


class Configuration:
    _default = None

    def __init__(self):
        # $TFOP_APISERVER is this stack's kubeconfig: the REST facade of a running operator
        self.host = os.environ.get("TFOP_APISERVER", "http://127.0.0.1:8443")
        self.temp_folder_path = None
        self.api_key = {}
        self.api_key_prefix = {}
        self.username = ""
        self.password = ""
        self.verify_ssl = True
        self.ssl_ca_cert = None
        self.cert_file = None
        self.key_file = None
        self.assert_hostname = None
        self.connection_pool_maxsize = 4
        self.proxy = None
        self.safe_chars_for_path_param = ""
        self.timeout = 120.0
        self.logger = {"package_logger": logging.getLogger("tf_operator_b200.sdk"),
                       "urllib3_logger": logging.getLogger("urllib3")}
        self.logger_format = "%(asctime)s %(levelname)s %(message)s"
        self.logger_file = None
        self._debug = False

    @classmethod
    def set_default(cls, default) -> None:
        cls._default = copy.copy(default)

    @classmethod
    def get_default_copy(cls) -> "Configuration":
        return copy.copy(cls._default) if cls._default is not None else Configuration()

    @property
    def debug(self) -> bool:
        return self._debug

    @debug.setter
    def debug(self, value: bool) -> None:
        self._debug = bool(value)
        for lg in self.logger.values():
            lg.setLevel(logging.DEBUG if value else logging.WARNING)

    def get_api_key_with_prefix(self, identifier: str):
        key = self.api_key.get(identifier)
        if key is None:
            return None
        prefix = self.api_key_prefix.get(identifier)
        return f"{prefix} {key}" if prefix else key

    def get_basic_auth_token(self) -> str:
        import base64
        return "Basic " + base64.b64encode(f"{self.username}:{self.password}".encode()).decode()

    def auth_settings(self) -> dict:
        return {}

    def to_debug_report(self) -> str:
        from ..version import VERSION
        return ("Python SDK Debug Report:\n"
                f"OS: {sys.platform}\nPython Version: {sys.version}\n"
                f"Version of the API: v1\nSDK Package Version: {VERSION}")
