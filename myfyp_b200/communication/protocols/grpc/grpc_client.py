"""gRPC client (reference: communication/protocols/grpc/grpc_client.py:37-206)."""

from __future__ import annotations

import random
import time
from typing import Any, Dict, List, Optional

from myfyp_b200.communication.protocols.client import Client
from myfyp_b200.communication.protocols.exceptions import CommunicationError, NeighborNotConnectedError, peer_looks_dead
from myfyp_b200.communication.protocols.grpc.grpc_neighbors import GrpcNeighbors
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings


class GrpcClient(Client):
    def __init__(self, self_addr: str, neighbors: GrpcNeighbors) -> None:
        self.__self_addr = self_addr
        self.__neighbors = neighbors

    def build_message(self, cmd: str, args: Optional[List[str]] = None, round: Optional[int] = None) -> Dict[str, Any]:
        args = [] if args is None else [str(a) for a in args]
        return {
            "source": self.__self_addr,
            "ttl": Settings.TTL,
            "hash": hash(str(cmd) + str(args) + str(time.time()) + str(random.randint(0, 100000))),
            "cmd": cmd,
            "args": args,
            "round": -1 if round is None else round,
        }

    def build_weights(self, cmd: str, round: int, serialized_model: Any, contributors: Optional[List[str]] = None, weight: int = 1) -> Dict[str, Any]:
        return {
            "source": self.__self_addr,
            "round": round,
            "weights": serialized_model,
            "contributors": [] if contributors is None else list(contributors),
            "weight": weight,
            "cmd": cmd,
        }

    def send(self, nei: str, msg: Dict[str, Any], create_connection: bool = False, raise_error: bool = False, remove_on_error: bool = True) -> None:
        channel = None
        try:
            try:
                stub = self.__neighbors.get(nei)[1]
            except KeyError as e:
                raise NeighborNotConnectedError(f"Neighbor {nei} not found.") from e
            if stub is None and create_connection:
                channel, stub = self.__neighbors.make_stub(nei)  # temporary channel to a non-direct neighbour
            if stub is None:
                raise NeighborNotConnectedError("Neighbor not directly connected (no stub and create_connection is false).")
            payload = msg
            if "weight" in msg and not stub.local and not isinstance(msg["weights"], (bytes, bytearray)):
                from myfyp_b200.parallel.ipc import encode_descriptor

                payload = dict(msg)
                payload["weights"] = encode_descriptor(msg["weights"])
            if stub.local and "weight" not in msg and Settings.DEDUP_BEFORE_SEND:
                srv = self.__neighbors.local_server(nei)  # in-process short-circuit of the flood: the receiver has seen this hash already
                if srv is not None and srv.gossiper.was_processed(msg["hash"]):
                    return
            if not raise_error and "weight" not in msg and hasattr(stub, "send_oneway"):
                err = stub.send_oneway(payload)  # the caller does not want the peer's verdict: do not wait for it
            else:
                err = stub.send(payload, Settings.GRPC_TIMEOUT)
            if err:
                raise CommunicationError(f"Error while sending a message: {msg['cmd']!r}: {err!r}")
        except Exception as e:  # noqa: BLE001
            logger.debug(self.__self_addr, f"Cannot send message {msg['cmd']!r} to {nei}. Error: {e}")
            if remove_on_error:
                direct = self.__neighbors.is_direct(nei)
                self.__neighbors.remove(nei, disconnect_msg=True)
                if direct and peer_looks_dead(e):
                    self.__neighbors.report_lost(nei)
            if raise_error:
                raise
        finally:
            if channel is not None:
                channel.close()

    def broadcast(self, msg: Dict[str, Any], node_list: Optional[List[str]] = None) -> None:
        nodes = node_list if node_list is not None else list(self.__neighbors.direct_addrs())
        if "hash" in msg and node_list is None:
            # flood bookkeeping: who has (or is about to get) this message - receivers do not forward to them.  (Every first-hop
            # receiver starts its own shared set from this list: handing them ONE set from here removes the flood's last
            # redundancy - each node is then targeted exactly once - and made federations stall now and then.)
            msg["seen"] = [self.__self_addr, *nodes]
        for n in nodes:
            self.send(n, msg)
