"""Minimal REST layer of the swagger runtime (reference: sdk/python/kubeflow/tfjob/rest.py -
``RESTClientObject`` over urllib3, ``ApiException``). urllib from the standard library; same
method names and the same exception shape."""
from __future__ import annotations

import json
import urllib.error
import urllib.parse
import urllib.request
from typing import Any, Dict, Optional


class ApiException(Exception):
    def __init__(self, status=None, reason=None, http_resp=None):
        if http_resp is not None:
            self.status, self.reason = http_resp.status, http_resp.reason
            self.body, self.headers = http_resp.data, http_resp.getheaders()
        else:
            self.status, self.reason, self.body, self.headers = status, reason, None, None
        super().__init__(str(self))

    def __str__(self) -> str:
        msg = f"({self.status})\nReason: {self.reason}\n"
        if self.headers:
            msg += f"HTTP response headers: {self.headers}\n"
        if self.body:
            msg += f"HTTP response body: {self.body}\n"
        return msg


class RESTResponse:
    def __init__(self, status: int, reason: str, data: bytes, headers: Dict[str, str]):
        self.status, self.reason, self.data, self._headers = status, reason, data, headers

    def getheaders(self) -> Dict[str, str]:
        return self._headers

    def getheader(self, name: str, default=None):
        return self._headers.get(name, default)


class RESTClientObject:
    def __init__(self, configuration, pools_size: int = 4, maxsize: Optional[int] = None):
        self.configuration = configuration

    def request(self, method: str, url: str, query_params=None, headers=None, body=None,
                post_params=None, _preload_content: bool = True, _request_timeout=None) -> RESTResponse:
        method = method.upper()
        assert method in ("GET", "HEAD", "DELETE", "POST", "PUT", "PATCH", "OPTIONS")
        if query_params:
            url += "?" + urllib.parse.urlencode(query_params)
        headers = dict(headers or {})
        data = None
        if body is not None:
            headers.setdefault("Content-Type", "application/json")
            data = body if isinstance(body, (bytes, str)) else json.dumps(body)
            if isinstance(data, str):
                data = data.encode()
        req = urllib.request.Request(url, data=data, method=method, headers=headers)
        timeout = _request_timeout or getattr(self.configuration, "timeout", 120.0)
        try:
            with urllib.request.urlopen(req, timeout=timeout) as r:
                resp = RESTResponse(r.status, r.reason, r.read(), dict(r.headers))
        except urllib.error.HTTPError as e:
            resp = RESTResponse(e.code, e.reason, e.read(), dict(e.headers))
            raise ApiException(http_resp=resp) from None
        except urllib.error.URLError as e:
            raise ApiException(status=0, reason=str(e.reason)) from None
        return resp

    def GET(self, url, **kw): return self.request("GET", url, **kw)
    def HEAD(self, url, **kw): return self.request("HEAD", url, **kw)
    def OPTIONS(self, url, **kw): return self.request("OPTIONS", url, **kw)
    def DELETE(self, url, **kw): return self.request("DELETE", url, **kw)
    def POST(self, url, **kw): return self.request("POST", url, **kw)
    def PUT(self, url, **kw): return self.request("PUT", url, **kw)
    def PATCH(self, url, **kw): return self.request("PATCH", url, **kw)


def to_json(obj: Any) -> str:
    return json.dumps(obj, default=str)
