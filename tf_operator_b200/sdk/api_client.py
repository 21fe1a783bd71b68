"""``ApiClient``: generic (de)serialising HTTP client of the swagger runtime (reference:
sdk/python/kubeflow/tfjob/api_client.py - sanitize_for_serialization, deserialize, call_api,
select_header_*). The job clients do not use it (neither do the reference's); it serves code that
wants raw REST access with the SDK models, e.g.

    api = ApiClient()
    job = api.call_api("/apis/kubeflow.org/v1/namespaces/{namespace}/tfjobs/{name}", "GET",
                       path_params={"namespace": "default", "name": "mnist"},
                       response_type="V1TFJob")
"""
from __future__ import annotations

import datetime
import json
import re
from typing import Any, Dict, Optional

from . import models
from .configuration import Configuration
from .rest import RESTClientObject

_PRIMITIVES = (float, bool, bytes, str, int)
_NATIVE = {"int": int, "float": float, "str": str, "bool": bool, "object": object}


class ApiClient:
    def __init__(self, configuration: Optional[Configuration] = None, header_name=None,
                 header_value=None, cookie=None):
        self.configuration = configuration or Configuration.get_default_copy()
        self.rest_client = RESTClientObject(self.configuration)
        self.default_headers: Dict[str, str] = {}
        if header_name is not None:
            self.default_headers[header_name] = header_value
        self.cookie = cookie
        self.user_agent = "tf_operator_b200-sdk/python"

    @property
    def user_agent(self) -> str:
        return self.default_headers["User-Agent"]

    @user_agent.setter
    def user_agent(self, value: str) -> None:
        self.default_headers["User-Agent"] = value

    def set_default_header(self, name: str, value: str) -> None:
        self.default_headers[name] = value

    # ---- serialisation ---------------------------------------------------------------------
    def sanitize_for_serialization(self, obj: Any) -> Any:
        """Model -> dict with JSON (camelCase) keys, recursively; None fields dropped."""
        if obj is None or isinstance(obj, _PRIMITIVES):
            return obj
        if isinstance(obj, (list, tuple)):
            return [self.sanitize_for_serialization(x) for x in obj]
        if isinstance(obj, (datetime.datetime, datetime.date)):
            return obj.isoformat()
        if isinstance(obj, models.Model):
            return models.to_wire(obj)
        if isinstance(obj, dict):
            return {k: self.sanitize_for_serialization(v) for k, v in obj.items() if v is not None}
        raise TypeError(f"cannot serialise {type(obj).__name__}")

    def deserialize(self, response, response_type: str) -> Any:
        data = response.data if hasattr(response, "data") else response
        if isinstance(data, (bytes, str)):
            try:
                data = json.loads(data)
            except ValueError:
                data = data.decode() if isinstance(data, bytes) else data
        return self._deserialize(data, response_type)

    def _deserialize(self, data: Any, klass) -> Any:
        if data is None:
            return None
        if isinstance(klass, str):
            m = re.match(r"list\[(.*)\]", klass)
            if m:
                return [self._deserialize(x, m.group(1)) for x in data]
            m = re.match(r"dict\(([^,]*), (.*)\)", klass)
            if m:
                return {k: self._deserialize(v, m.group(2)) for k, v in data.items()}
            if klass in _NATIVE:
                klass = _NATIVE[klass]
            elif klass == "datetime":
                return datetime.datetime.fromisoformat(str(data).replace("Z", "+00:00"))
            else:
                klass = getattr(models, klass)
        if klass is object:
            return data
        if klass in (int, float, str, bool):
            return klass(data)
        if isinstance(klass, type) and issubclass(klass, models.Model):
            kwargs = {}
            for attr, typ in klass.swagger_types.items():
                key = klass.attribute_map[attr]
                if isinstance(data, dict) and key in data:
                    kwargs[attr] = self._deserialize(data[key], typ)
            return klass(**kwargs)
        return data

    # ---- calls -----------------------------------------------------------------------------
    def call_api(self, resource_path: str, method: str, path_params=None, query_params=None,
                 header_params=None, body=None, response_type: Optional[str] = None,
                 _return_http_data_only: bool = True, _preload_content: bool = True,
                 _request_timeout=None, **_ignored):
        headers = dict(self.default_headers)
        headers.update(header_params or {})
        if self.cookie:
            headers["Cookie"] = self.cookie
        for k, v in (path_params or {}).items():
            resource_path = resource_path.replace("{%s}" % k, str(v))
        url = self.configuration.host.rstrip("/") + resource_path
        resp = self.rest_client.request(method, url, query_params=query_params, headers=headers,
                                        body=self.sanitize_for_serialization(body),
                                        _request_timeout=_request_timeout)
        if not _preload_content:
            return resp
        data = self.deserialize(resp, response_type) if response_type else (
            json.loads(resp.data) if resp.data else None)
        return data if _return_http_data_only else (data, resp.status, resp.getheaders())

    @staticmethod
    def select_header_accept(accepts):
        if not accepts:
            return None
        accepts = [a.lower() for a in accepts]
        return "application/json" if "application/json" in accepts else ", ".join(accepts)

    @staticmethod
    def select_header_content_type(content_types):
        if not content_types:
            return "application/json"
        content_types = [c.lower() for c in content_types]
        return "application/json" if ("application/json" in content_types or "*/*" in content_types) \
            else content_types[0]
