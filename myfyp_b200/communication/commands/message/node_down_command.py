"""``node_down``: "I found my direct neighbour <addr> dead" - failure dissemination.  No reference counterpart.

In the reference (and here until round 2) every node finds out about a death on its own: direct neighbours when a send fails or the
beats stop, everybody else only when the dead node's flooded beats have been missing for ``HEARTBEAT_TIMEOUT``
(p2pfl/communication/protocols/heartbeater.py:80-113).  A vote needs every known node's ballot, so one death stalled the whole
federation for that time-out (64 nodes, 3 s time-out: 3 s of a 7 s experiment).  The first node that notices now floods this
message, and a node that stops in an orderly way floods it about itself before it goes; receivers that know the dead node only through the flood drop it at once (and ignore its still-travelling messages for
two beat periods), receivers that are its direct neighbours keep trusting their own link.  A node declared dead by mistake comes
back with its next beat.
"""

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.management.logger import logger
from myfyp_b200.settings import Settings


class NodeDownCommand(Command):
    def __init__(self, self_addr: str, neighbors) -> None:
        self.__addr, self.__neighbors = self_addr, neighbors

    @staticmethod
    def get_name() -> str:
        return "node_down"

    def execute(self, source: str, round: int, *args, **kwargs) -> None:
        if not args or not Settings.FAILURE_DISSEMINATION:
            return
        dead = str(args[0])
        if dead == self.__addr:
            return  # (dead == source: the node's own goodbye, flooded by ComposedProtocol.stop())
        if self.__neighbors.bury(dead, 2.0 * Settings.HEARTBEAT_PERIOD + 0.2):
            logger.info(self.__addr, f"👋 {dead} said goodbye: removed." if source == dead else f"☠️ {source} reports {dead} dead: removed.")
