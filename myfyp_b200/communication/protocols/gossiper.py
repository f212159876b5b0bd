"""Gossiper: TTL-flood sender thread + the synchronous model-gossip loop.

Reference: communication/protocols/gossiper.py:31-239.  Behavioural contract kept:
dedupe ring of ``AMOUNT_LAST_MESSAGES_SAVED`` hashes, at most
``GOSSIP_MESSAGES_PER_PERIOD`` sends per ``GOSSIP_PERIOD``, model gossip to
``min(GOSSIP_MODELS_PER_ROUND, n)`` random candidates per tick, exit when no candidate
is left or the status did not change for ``GOSSIP_EXIT_ON_X_EQUAL_ROUNDS`` ticks.

Deliberate differences:
* the sender thread sleeps on a condition variable, so an idle gossiper costs nothing
  and ``GOSSIP_PERIOD = 0`` does not busy-spin;
* periods are *minimum spacings*: a tick that changed the status is followed
  immediately by the next one, an unchanged one waits (interruptibly) for the period;
* "X equal rounds" means what the docs say: the last X recorded statuses are all equal.
"""

from __future__ import annotations

import collections
from concurrent.futures import ThreadPoolExecutor
import random
import threading
import time
from typing import Any, Callable, Deque, List, Optional, Tuple

from myfyp_b200.communication.protocols.client import Client
from myfyp_b200.management.logger import logger
from myfyp_b200.management.trace import trace
from myfyp_b200.settings import Settings


_POOL: Optional[ThreadPoolExecutor] = None
_POOL_LOCK = threading.Lock()


def _send_pool() -> ThreadPoolExecutor:
    global _POOL
    if _POOL is None:
        with _POOL_LOCK:
            if _POOL is None:
                _POOL = ThreadPoolExecutor(max_workers=max(4, 4 * Settings.GOSSIP_SEND_WORKERS), thread_name_prefix="gossip-send")
    return _POOL


_inline = threading.local()  # forwarding depth of the current thread (add_message, inline_ok)


class Gossiper(threading.Thread):
    def __init__(self, self_addr: str, client: Client, period: Optional[float] = None, messages_per_period: Optional[int] = None) -> None:
        super().__init__(daemon=True, name=f"gossiper-thread-{self_addr}")
        self.__self_addr = self_addr
        self.__client = client
        self.period = period
        self.messages_per_period = messages_per_period
        self.__processed: "collections.OrderedDict[int, None]" = collections.OrderedDict()
        self.__processed_lock = threading.Lock()
        self.__pending: Deque[Tuple[Any, List[str], float]] = collections.deque()  # (message, recipients, not before [monotonic])
        self.__cv = threading.Condition()
        self.__terminate = threading.Event()
        self.__pool: Optional[ThreadPoolExecutor] = None

    # ------------------------------------------------------------------ thread control
    def start(self) -> None:
        logger.debug(self.__self_addr, "🏁 Starting gossiper...")
        super().start()

    def stop(self) -> None:
        logger.debug(self.__self_addr, "🛑 Stopping gossiper...")
        self.__terminate.set()
        with self.__cv:
            self.__cv.notify_all()
        if self.__pool is not None:  # (kept for subclasses that install a private pool)
            self.__pool.shutdown(wait=False)

    # ------------------------------------------------------------------ flood
    def add_message(self, msg: Any, pending_neis: List[str], coalesce_key: Any = None, linger: float = 0.0, inline_ok: bool = False) -> None:
        """Queue ``msg`` for ``pending_neis``.  With a ``coalesce_key`` a still-unsent message queued under the same key is
        *replaced* (its remaining recipients get the newer message instead): a burst of state announcements costs one
        message per neighbour, not one per state change.  ``linger`` holds a coalescable message back for that many
        seconds so that a burst really does collapse (the first deadline of a key is kept: nothing is starved);
        ``linger = 0`` with an entry already queued under the key sends the replacement at once."""
        if not pending_neis:
            return
        if inline_ok and coalesce_key is None and linger <= 0 and Settings.GOSSIP_INLINE_FORWARD:
            # (the caller vouches that ``pending_neis`` live in this process)
            # Forwarding a flooded message between nodes of ONE process: the next hop's handler is a function call, so the hop is
            # made on the thread that is delivering the message instead of waking this node's gossiper thread for it (at 64 nodes
            # a broadcast cost ~60 thread wake-ups under one interpreter lock).  Depth-limited: a long chain of hops falls back
            # to the queue, and so does every message a HANDLER emits (those go through broadcast_async, never through here).
            depth = getattr(_inline, "depth", 0)
            if depth < Settings.GOSSIP_INLINE_DEPTH:
                _inline.depth = depth + 1
                try:
                    for nei in pending_neis:
                        if self.__terminate.is_set():
                            return
                        self.__client.send(nei, msg)
                finally:
                    _inline.depth = depth
                return
        now = time.monotonic()
        with self.__cv:
            if coalesce_key is not None:
                for i, (old, neis, due) in enumerate(self.__pending):
                    if isinstance(old, dict) and old.get("_ckey") == coalesce_key:
                        merged = list(dict.fromkeys([*neis, *pending_neis]))
                        tagged = dict(msg)
                        tagged["_ckey"] = coalesce_key
                        self.__pending[i] = (tagged, merged, min(due, now + linger))
                        self.__cv.notify()
                        return
                msg = dict(msg)
                msg["_ckey"] = coalesce_key
            self.__pending.append((msg, list(pending_neis), now + (linger if coalesce_key is not None else 0.0)))
            self.__cv.notify()

    def check_and_set_processed(self, msg_hash: int) -> bool:
        """True the first time a hash is seen (and remembers it)."""
        with self.__processed_lock:
            if msg_hash in self.__processed:
                return False
            while len(self.__processed) > Settings.AMOUNT_LAST_MESSAGES_SAVED:
                self.__processed.popitem(last=False)
            self.__processed[msg_hash] = None
            return True

    def was_processed(self, msg_hash: int) -> bool:
        """Non-mutating probe used to skip redundant in-process sends."""
        return msg_hash in self.__processed

    def run(self) -> None:
        while not self.__terminate.is_set():
            period = Settings.GOSSIP_PERIOD if self.period is None else self.period
            budget = Settings.GOSSIP_MESSAGES_PER_PERIOD if self.messages_per_period is None else self.messages_per_period
            batch: List[Tuple[Any, List[str]]] = []
            with self.__cv:
                while not self.__terminate.is_set():
                    now = time.monotonic()
                    if any(due <= now for _, _, due in self.__pending):
                        break
                    nxt = min((due for _, _, due in self.__pending), default=None)
                    self.__cv.wait(timeout=1.0 if nxt is None else max(0.0, nxt - now))
                if self.__terminate.is_set():
                    return
                now = time.monotonic()
                kept: Deque[Tuple[Any, List[str], float]] = collections.deque()
                while self.__pending:
                    msg, neis, due = self.__pending.popleft()
                    if due > now or budget <= 0:  # lingering announcement, or this period's budget is spent
                        kept.append((msg, neis, due))
                    elif len(neis) <= budget:
                        batch.append((msg, neis))
                        budget -= len(neis)
                    else:
                        batch.append((msg, neis[:budget]))
                        kept.append((msg, neis[budget:], due))
                        budget = 0
                self.__pending = kept
            t0 = time.monotonic()
            for msg, neis in batch:
                if isinstance(msg, dict) and "_ckey" in msg:
                    # the coalescing tag is queue-local: on the wire it would make the RECEIVER's gossiper take this (forwarded)
                    # message for one of its own and replace it with its own next announcement
                    msg = {k: v for k, v in msg.items() if k != "_ckey"}
                for nei in neis:
                    if self.__terminate.is_set():
                        return
                    trace(self.__self_addr, f"flood {msg.get('cmd') if isinstance(msg, dict) else '?'}>{nei}")
                    self.__client.send(nei, msg)
            # only a *saturated* period needs pacing; otherwise go back to waiting
            if budget == 0 and period > 0:
                left = period - (time.monotonic() - t0)
                if left > 0:
                    self.__terminate.wait(left)

    # ------------------------------------------------------------------ synchronous model gossip
    def gossip_weights(
        self,
        early_stopping_fn: Callable[[], bool],
        get_candidates_fn: Callable[[], List[str]],
        status_fn: Callable[[], Any],
        model_fn: Callable[[str], Any],
        period: float,
        create_connection: bool,
        wait_fn: Optional[Callable[[float], None]] = None,
    ) -> None:
        # "X equal rounds" is a statement about TIME in the reference (its rounds are one period apart, gossiper.py:205-222).  The
        # loop here is also woken by state changes that have nothing to do with the status (ballots for the next round, other
        # nodes' announcements): such wake-ups re-evaluate the candidates but must not count, or ten of them within milliseconds
        # end the gossip while a peer still lacks models.  A tick is counted once per period without a status change.
        x = max(1, Settings.GOSSIP_EXIT_ON_X_EQUAL_ROUNDS)
        tick = period if period > 0 else 0.02
        last_seen: Optional[str] = None
        equal_ticks = 0
        tick_at = time.monotonic()
        while True:
            t0 = time.monotonic()
            if early_stopping_fn():
                logger.info(self.__self_addr, "Stopping model gossip process.")
                return
            neis = get_candidates_fn()
            if not neis:
                logger.info(self.__self_addr, "🤫 Gossip finished.")
                return
            logger.debug(self.__self_addr, f"👥 Gossip remaining nodes: {neis}")
            before = repr(status_fn())
            if before != last_seen:
                last_seen, equal_ticks, tick_at = before, 0, t0
            elif t0 - tick_at >= tick:
                equal_ticks += 1
                tick_at = t0
            if equal_ticks >= x:
                logger.info(self.__self_addr, f"⏹️  Gossiping exited for {x} equal rounds.")
                return
            outbox = []
            for nei in random.sample(neis, min(Settings.GOSSIP_MODELS_PER_ROUND, len(neis))):
                msg = model_fn(nei)
                if msg is not None:
                    outbox.append((nei, msg))

            def deliver(item) -> None:
                nei, msg = item
                logger.debug(self.__self_addr, f"🗣️ Gossiping model to {nei}.")
                trace(self.__self_addr, f"send_model>{nei}")
                self.__client.send(nei, msg, create_connection=create_connection)
                trace(self.__self_addr, f"sent_model>{nei}")

            if len(outbox) > 1 and Settings.GOSSIP_SEND_WORKERS > 1 and not getattr(self.__client, "in_process", False):
                # each send is a blocking request -> response (the peer pulls and files the model before it answers):
                # to several peers they overlap instead of adding up (the reference sends one after the other).
                # ONE pool per process: a pool per node meant hundreds of mostly idle threads at 64 nodes, and every submit of
                # every pool serialises on the executor module's global lock.  In-process transports run the handler on the
                # caller's thread anyway (nothing blocks on I/O): they deliver inline.
                list(_send_pool().map(deliver, outbox))
            else:
                for item in outbox:
                    deliver(item)
            # progress => go again at once; no progress => wait (interruptibly) for the period
            if repr(status_fn()) != before:
                continue
            left = period - (time.monotonic() - t0)
            if left > 0:
                if wait_fn is not None:
                    wait_fn(left)
                else:
                    time.sleep(left)
