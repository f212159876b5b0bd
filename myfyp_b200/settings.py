"""Global configuration knobs.

Same names and defaults as the reference's ``Settings`` class attributes
(reference: p2pfl/settings.py:34-150) so experiment scripts port unchanged.
Differences, on purpose:

* ``WAIT_HEARTBEATS_CONVERGENCE`` is a *property-like* value: when left at
  ``None`` it is derived from the current ``HEARTBEAT_TIMEOUT`` at read time
  (the reference freezes it at import).
* Extra B200 data-plane knobs live here too (``DEVICE_AGGREGATION``,
  ``NODES_PER_GPU`` ...), read at call time like everything else.
"""

from __future__ import annotations

import os

_HERE = os.path.dirname(os.path.abspath(__file__))


class _SettingsMeta(type):
    """Lets ``Settings.WAIT_HEARTBEATS_CONVERGENCE`` follow HEARTBEAT_TIMEOUT."""

    @property
    def WAIT_HEARTBEATS_CONVERGENCE(cls) -> float:  # noqa: N802
        v = cls.__dict__.get("_WAIT_HEARTBEATS_CONVERGENCE")
        return 0.2 * cls.HEARTBEAT_TIMEOUT if v is None else v

    @WAIT_HEARTBEATS_CONVERGENCE.setter
    def WAIT_HEARTBEATS_CONVERGENCE(cls, value) -> None:  # noqa: N802
        cls._WAIT_HEARTBEATS_CONVERGENCE = value


class Settings(metaclass=_SettingsMeta):
    """Mutable class-attribute settings bag (read at call time everywhere)."""

    # --- general -----------------------------------------------------------
    GRPC_TIMEOUT: float = 10
    LOG_LEVEL: str = "INFO"
    LOG_DIR: str = "logs"
    EXCLUDE_BEAT_LOGS: bool = True
    DISABLE_RAY: bool = False  # kept for script compatibility; Ray is not used
    FILE_LOGGING: bool = False  # the reference always writes logs/p2pfl.log; opt-in here
    SEED: int | None = None

    # --- heartbeat ---------------------------------------------------------
    HEARTBEAT_PERIOD: float = 2
    HEARTBEAT_TIMEOUT: float = 5

    # --- gossip (control messages) ------------------------------------------
    GOSSIP_PERIOD: float = 0.1
    TTL: int = 10
    GOSSIP_MESSAGES_PER_PERIOD: int = 100
    AMOUNT_LAST_MESSAGES_SAVED: int = 100

    # --- gossip (models) -----------------------------------------------------
    GOSSIP_MODELS_PERIOD: float = 1
    # An intermediate "models_aggregated" announcement (the node holds some, not yet all, models of the round) is held back this
    # long so that a burst of arrivals is announced once; the announcement of a COMPLETE set is never delayed.  (B200 addition.)
    ANNOUNCE_LINGER: float = 0.005
    # Model offers of one gossip tick are delivered by this many threads at once (1 = one after the other, like the reference).
    GOSSIP_SEND_WORKERS: int = 8
    # A node forwards models it did not train itself to a peer only after holding them this long: in a dense topology the origin
    # delivers its own model to that peer directly within milliseconds, and an early forward is a duplicate the peer must
    # refuse.  Own models are always offered at once.  (B200 addition; the reference forwards immediately, every second.)
    FORWARD_GRACE: float = 0.02
    GOSSIP_MODELS_PER_ROUND: int = 2
    GOSSIP_EXIT_ON_X_EQUAL_ROUNDS: int = 10
    FAILURE_DISSEMINATION: bool = True   # flood ``node_down`` when a direct neighbour is found dead (nodes further away need not wait for HEARTBEAT_TIMEOUT)
    GOSSIP_INLINE_FORWARD: bool = True   # in-process protocols: forward a flooded message on the delivering thread (no gossiper wake-up per hop)
    GOSSIP_INLINE_DEPTH: int = 12        # ... for at most this many nested hops; deeper ones are queued

    # --- TLS -------------------------------------------------------------------
    USE_SSL: bool = True
    CA_CRT: str = f"{_HERE}/certificates/ca.crt"
    SERVER_CRT: str = f"{_HERE}/certificates/server.crt"
    CLIENT_CRT: str = f"{_HERE}/certificates/client.crt"
    SERVER_KEY: str = f"{_HERE}/certificates/server.key"
    CLIENT_KEY: str = f"{_HERE}/certificates/client.key"

    # --- learning ----------------------------------------------------------------
    TRAIN_SET_SIZE: int = 4
    VOTE_TIMEOUT: float = 60
    AGGREGATION_TIMEOUT: float = 300
    _WAIT_HEARTBEATS_CONVERGENCE = None
    RESOURCE_MONITOR_PERIOD: float = 1

    # --- B200 data plane (no counterpart in the reference) -------------------------
    # Weights messages between nodes that can reach each other's HBM carry a slab
    # descriptor instead of pickled bytes; aggregation then pulls straight from the
    # sender's slab inside the aggregation kernel.
    DEVICE_WEIGHTS: bool = True
    # Event-driven control plane: waits are woken by condition variables; the
    # periods above become *minimum spacings* rather than fixed sleeps.
    EVENT_DRIVEN: bool = True
    # Short-circuit redundant flood sends when the receiver lives in this process
    # and has already seen the hash.
    DEDUP_BEFORE_SEND: bool = True
    # Malformed model from a peer: "stop" (reference behaviour) or "drop" the
    # message and evict the sender.
    ON_BAD_MODEL: str = "drop"
    # Let nodes (re)join a running experiment.
    ALLOW_REJOIN: bool = True
    # gRPC peers that live in this process are called directly (no socket, descriptors stay objects).
    GRPC_LOCAL_SHORTCUT: bool = True
    # Default experiment name (reference hard-codes "experiment").
    EXPERIMENT_NAME: str = "experiment"
    # Directory for per-round checkpoints (wire-layout blob + JSON side-car per node); None = off.  (B200 addition; the
    # reference has no disk checkpoints.)
    CHECKPOINT_DIR: str | None = None

    # --- loading (the reference has no files, environment variables or validation) ------------------------
    @classmethod
    def names(cls) -> list:
        """Every setting (upper-case attribute), in definition order."""
        return [k for k in vars(cls) if k.isupper() and not k.startswith("_")] + ["WAIT_HEARTBEATS_CONVERGENCE"]

    @classmethod
    def update(cls, values: dict, strict: bool = True) -> None:
        """Set several settings at once; a value is coerced to the type of the current one (``"0.5"`` -> 0.5, ``"false"`` ->
        False).  Unknown names raise ``KeyError`` (``strict``) - a typo must not silently configure nothing."""
        known = set(cls.names())
        for key, value in values.items():
            if key not in known:
                if strict:
                    raise KeyError(f"unknown setting {key!r}")
                continue
            cur = getattr(cls, key)
            if isinstance(value, str) and cur is not None and not isinstance(cur, str):
                if isinstance(cur, bool):
                    value = value.strip().lower() in ("1", "true", "yes", "on")
                else:
                    value = type(cur)(float(value)) if isinstance(cur, int) else type(cur)(value)
            setattr(cls, key, value)

    @classmethod
    def update_from_env(cls, prefix: str = "MYFYP_B200_") -> list:
        """``MYFYP_B200_TRAIN_SET_SIZE=8`` etc.; returns the names that were set."""
        found = {k[len(prefix):]: v for k, v in os.environ.items() if k.startswith(prefix)}
        cls.update(found, strict=False)
        return sorted(k for k in found if k in set(cls.names()))

    @classmethod
    def load_file(cls, path: str) -> None:
        """A JSON or YAML mapping of setting names to values."""
        with open(path) as f:
            text = f.read()
        if path.endswith((".yaml", ".yml")):
            import yaml

            data = yaml.safe_load(text)
        else:
            import json

            data = json.loads(text)
        if not isinstance(data, dict):
            raise ValueError(f"{path}: expected a mapping of setting names to values")
        cls.update(data)
