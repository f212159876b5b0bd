"""REST client for "P2PFL Web Services" (reference: p2pfl/management/p2pfl_web_services.py:58-270).

Endpoints: ``POST /node``, ``/node-log``, ``/node-metric/local|global|system`` with an
``x-api-key`` header.  Failures never propagate into the learning path.
"""

from __future__ import annotations

import datetime
import logging
from typing import Dict


class P2pflWebServicesError(Exception):
    def __init__(self, code: int, message: str) -> None:
        super().__init__(f"Error {code}: {message}")
        self.code, self.message = code, message


class P2pflWebServices:
    def __init__(self, url: str, key: str) -> None:
        self._url = url if url.endswith("/") else url + "/"
        self._key = key
        self.node_id: Dict[str, int] = {}

    def _headers(self) -> Dict[str, str]:
        return {"Content-Type": "application/json", "x-api-key": self._key}

    def _post(self, path: str, data: dict) -> dict:
        import requests

        r = requests.post(self._url + path, json=data, headers=self._headers(), timeout=5)
        if r.status_code not in (200, 201):
            raise P2pflWebServicesError(r.status_code, r.text)
        try:
            return r.json()
        except Exception:
            return {}

    def register_node(self, node: str, is_simulated: bool) -> None:
        res = self._post("node", {"address": node, "is_simulated": is_simulated})
        self.node_id[node] = res.get("node_id", -1)

    def unregister_node(self, node: str) -> None:
        self.node_id.pop(node, None)

    def send_log(self, time: datetime.datetime, node: str, level: int, message: str) -> None:
        if node in self.node_id:
            self._post("node-log", {"time": str(time), "node_id": self.node_id[node], "level": level, "message": message})

    def send_local_metric(self, exp: str, round: int, metric: str, node: str, value: float, step: int) -> None:
        if node in self.node_id:
            self._post(
                "node-metric/local",
                {"node_id": self.node_id[node], "exp": exp, "round": round, "metric_name": metric, "metric_value": value, "step": step},
            )

    def send_global_metric(self, exp: str, round: int, metric: str, node: str, value: float) -> None:
        if node in self.node_id:
            self._post(
                "node-metric/global",
                {"node_id": self.node_id[node], "exp": exp, "round": round, "metric_name": metric, "metric_value": value},
            )

    def send_system_metric(self, node: str, metric: str, value: float, time: datetime.datetime) -> None:
        if node in self.node_id:
            self._post(
                "node-metric/system",
                {"node_id": self.node_id[node], "metric_name": metric, "metric_value": value, "time": str(time)},
            )

    def get_pending_actions(self) -> None:
        raise NotImplementedError("not offered by the web service (neither in the reference)")


class WebLogHandler(logging.Handler):
    """Forward text logs of registered nodes to the web service."""

    def __init__(self, web: P2pflWebServices) -> None:
        super().__init__()
        self._web = web

    def emit(self, record: logging.LogRecord) -> None:
        node = getattr(record, "node", None)
        if node is None:
            return
        try:
            self._web.send_log(datetime.datetime.fromtimestamp(record.created), node, record.levelno, record.getMessage())
        except Exception:
            pass
