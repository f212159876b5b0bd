"""Weight diffusion shared by the stages that push a model to peers (initial model, aggregated model).

A diffusion is "keep offering ``command`` to every peer in ``pending()`` until nobody is pending, the experiment stops, or
the picture stops changing"; the loop itself is the protocol's ``gossip_weights`` (event driven, see ``Gossiper``).  What is
offered is decided per peer: a device descriptor when the peer can dereference it (same process, or an IPC arena on the same
box), the reference's pickle otherwise - built at most once per kind and diffusion, not once per peer.
"""

from __future__ import annotations

from typing import Any, Callable, Dict, List


def transport_payload(model, communication_protocol, nei: str):
    """Device descriptor when ``nei`` can dereference it, the reference pickle otherwise."""
    enc = getattr(model, "encode_for_transport", None)
    if enc is None:
        return model.encode_parameters()
    return enc(communication_protocol.fabric_reachable(nei))


class Diffusion:
    def __init__(self, state, communication_protocol, command: str, pending: Callable[[], List[str]], model: Callable[[], Any], cache_payload: bool = True) -> None:
        self.state, self.protocol, self.command, self.pending, self.model = state, communication_protocol, command, pending, model
        self._payloads: Dict[bool, Any] = {}
        self._cache = cache_payload

    def _message(self, nei: str) -> Any:
        if self.state.round is None:
            raise Exception("Round not initialized.")
        reach = bool(self.protocol.fabric_reachable(nei))
        if not self._cache or reach not in self._payloads:
            self._payloads[reach] = transport_payload(self.model(), self.protocol, nei)
        return self.protocol.build_weights(self.command, self.state.round, self._payloads[reach])

    def run(self) -> None:
        self.protocol.gossip_weights(lambda: self.state.round is None, self.pending, self.pending, self._message, wait_fn=self.state.wait_change)
