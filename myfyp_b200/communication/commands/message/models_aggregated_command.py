"""``models_aggregated`` (reference: commands/message/models_agregated_command.py:26-56)."""

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.management.logger import logger
from myfyp_b200.management.trace import trace


class ModelsAggregatedCommand(Command):
    def __init__(self, state) -> None:
        self.state = state

    @staticmethod
    def get_name() -> str:
        return "models_aggregated"

    def execute(self, source: str, round: int, *args, **kwargs) -> None:
        trace(self.state.addr, f"announce<{source} {len(args)}")
        if round == self.state.round:
            # Within a round a node's held set only grows, so announcements are merged, not overwritten: two
            # announcements of one node can overtake each other (concurrent handlers, async fan-out), and a stale
            # "last writer" would make every later FedAvg offer overlap what the node already holds -> rejected
            # forever -> the round ends by timeout.  (The reference overwrites, models_agregated_command.py:49-51.)
            cur = self.state.models_aggregated.get(source, [])
            self.state.models_aggregated[source] = cur + [a for a in args if a not in cur]
            # read by the partial-model gossip of train-set members only (TrainStage): nobody else is woken
            ts = self.state.train_set
            if not ts or self.state.addr in ts:
                self.state.notify()
        elif self.state.round is not None and round == self.state.round + 1:
            # the sender is already one round ahead of us: keep its announcement for when we get there
            nxt = self.state.models_aggregated_next.setdefault(source, [])
            nxt += [a for a in args if a not in nxt]
        else:
            logger.debug(self.state.addr, f"models_aggregated from {source} in a late round ({round} != {self.state.round}). Ignored.")
