"""``vote_train_set`` (reference: commands/message/vote_train_set_command.py:28-74)."""

import contextlib

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.management.logger import logger


class VoteTrainSetCommand(Command):
    def __init__(self, state) -> None:
        self.state = state

    @staticmethod
    def get_name() -> str:
        return "vote_train_set"

    def execute(self, source: str, round: int, *args, **kwargs) -> None:
        st = self.state
        if st.round is None:
            logger.debug(st.addr, f"Vote received from {source} while not learning. Ignored.")
            return
        # Votes may arrive one round early (the sender finished its round first).  They are kept apart from the current round's
        # ballots: a node still waiting in THIS round's vote (e.g. for the ballot of a peer that just died) would otherwise see a
        # fast peer's next-round ballot overwrite its current one, tally different ballots than everybody else and end up with
        # a different train set - after which every round waits for time-outs.  (The reference keeps one dict,
        # vote_train_set_command.py:60-71, and has this failure mode.)
        if round in (st.round, st.round + 1):
            votes = list(args)
            tmp = {}
            for i in range(0, len(votes) - 1, 2):
                tmp[votes[i]] = int(votes[i + 1])
            with st.train_set_votes_lock:
                if round == st.round:
                    st.train_set_votes[source] = tmp
                else:
                    st.train_set_votes_next[source] = tmp
                have = len(st.train_set_votes)
            with contextlib.suppress(Exception):
                st.wait_votes_ready_lock.release()
            # the vote stage says how many ballots it is waiting for: waking it for each of the 63 that do not complete the box is
            # 63 thread hand-offs under the interpreter lock per node and round (a neighbour change or a stop wakes it anyway)
            need = st.votes_needed
            if need is None or round != st.round or have >= need:
                st.notify()
        else:
            logger.debug(st.addr, f"Vote from {source} in a late round. Ignored. {round} != {st.round} / {st.round + 1}")
