"""Address parsing (reference: communication/protocols/grpc/address.py:26-99): ``ip:port``, ``ip`` (random free
port), ``[v6]:port``, ``unix:///abs/path``."""

from __future__ import annotations

import os
import socket
from ipaddress import ip_address
from typing import Optional


class AddressParser:
    def __init__(self, address: str) -> None:
        self.host: Optional[str] = None
        self.port: Optional[int] = None
        self.is_v6: Optional[bool] = None
        self.unix_domain = False
        if address.startswith("unix://") and os.path.isabs(address[len("unix://") :]):
            self.unix_domain, self.host = True, address
            return
        try:
            raw_host, _, raw_port = address.rpartition(":")
            if raw_host != "" and not (address.count(":") > 1 and not address.startswith("[")):
                self.port = int(raw_port)
                if not 1 <= self.port <= 65535:
                    raise ValueError("Port number is invalid.")
            else:
                raw_host = address
                with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
                    s.bind(("", 0))
                    self.port = s.getsockname()[1]
            self.host = raw_host.translate({ord(c): None for c in "[]"})
            self.is_v6 = ip_address(self.host).version == 6
        except ValueError:
            self.host = self.port = self.is_v6 = None

    def get_parsed_address(self) -> str:
        if self.unix_domain:
            return self.host  # type: ignore[return-value]
        if self.host is None:
            raise ValueError("The address is invalid.")
        return f"[{self.host}]:{self.port}" if self.is_v6 else f"{self.host}:{self.port}"
