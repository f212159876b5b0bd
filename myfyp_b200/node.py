"""``Node``: the user-facing facade that wires protocol + aggregator + learner + state + workflow.

Public API = the reference's (p2pfl/node.py:89-413): ``Node(model, data, address, learner, aggregator,
protocol, simulation)``, ``start/stop/connect/disconnect/get_neighbors``, ``set_start_learning(rounds, epochs)``,
``set_stop_learning``, ``set_model/set_data/set_learner/get_model/get_data``; attributes ``addr``, ``state``,
``learner``, ``aggregator``, ``learning_workflow``.

Differences: the default protocol is the in-memory one when the address is not an IP (gRPC otherwise, as in
the reference); the model is placed on a GPU by the :class:`~myfyp_b200.parallel.fabric.Fabric`; a node that
connects to a *learning* network is brought into the running experiment (rejoin) - the reference
explicitly does not support that (node.py:144-145).
"""

from __future__ import annotations

import contextlib
import threading
import traceback
from typing import Any, Dict, Optional, Type

from myfyp_b200.communication.commands.message.metrics_command import MetricsCommand
from myfyp_b200.communication.commands.message.model_initialized_command import ModelInitializedCommand
from myfyp_b200.communication.commands.message.models_aggregated_command import ModelsAggregatedCommand
from myfyp_b200.communication.commands.message.models_ready_command import ModelsReadyCommand
from myfyp_b200.communication.commands.message.start_learning_command import StartLearningCommand
from myfyp_b200.communication.commands.message.stop_learning_command import StopLearningCommand
from myfyp_b200.communication.commands.message.vote_train_set_command import VoteTrainSetCommand
from myfyp_b200.communication.commands.weights.full_model_command import FullModelCommand
from myfyp_b200.communication.commands.weights.init_model_command import InitModelCommand
from myfyp_b200.communication.commands.weights.partial_model_command import PartialModelCommand
from myfyp_b200.communication.protocols.communication_protocol import CommunicationProtocol
from myfyp_b200.exceptions import LearnerRunningException, NodeRunningException, ZeroRoundsException
from myfyp_b200.learning.aggregators.aggregator import Aggregator
from myfyp_b200.learning.aggregators.fedavg import FedAvg
from myfyp_b200.learning.learner import Learner, LearnerFactory
from myfyp_b200.learning.p2pfl_model import P2PFLModel
from myfyp_b200.management.logger import logger
from myfyp_b200.node_state import NodeState
from myfyp_b200.settings import Settings
from myfyp_b200.stages.workflows import LearningWorkflow


def _default_protocol(address: str) -> Type[CommunicationProtocol]:
    from myfyp_b200.communication.protocols.grpc.grpc_communication_protocol import GrpcCommunicationProtocol

    return GrpcCommunicationProtocol


class Node:
    def __init__(
        self,
        model: P2PFLModel,
        data,
        address: str = "127.0.0.1",
        learner: Optional[Type[Learner]] = None,
        aggregator: Optional[Aggregator] = None,
        protocol: Optional[Type[CommunicationProtocol]] = None,
        simulation: bool = False,
        device: Any = None,
        **kwargs,
    ) -> None:
        if protocol is None:
            protocol = _default_protocol(address)
        self._communication_protocol = protocol(address)
        self.addr = self._communication_protocol.get_address()

        self.aggregator = FedAvg() if aggregator is None else aggregator
        self.aggregator.node_name = self.addr

        # place the model on the fabric (round-robin over the visible GPUs; CPU without a GPU)
        if hasattr(model, "to") and hasattr(model, "slab"):
            from myfyp_b200.parallel.fabric import Fabric

            dev = device if device is not None else (model.device if str(model.device) != "cpu" else Fabric.get().place())
            model.to(dev)

        if learner is None:
            learner = LearnerFactory.create_learner(model)
        self.learner: Learner = learner(model, data, self.addr, self.aggregator, **kwargs) if isinstance(learner, type) else learner
        if not isinstance(learner, type):
            self.learner.set_addr(self.addr)
        if simulation:
            from myfyp_b200.learning.simulation import try_init_learner_with_simulation

            self.learner = try_init_learner_with_simulation(self.learner)

        self.__running = False
        self.state = NodeState(self.addr, simulation=simulation)
        self.state.is_direct = getattr(getattr(self._communication_protocol, "_neighbors", None), "is_direct", None)  # commands wake waiting stages selectively
        self.simulation = simulation
        self.learning_workflow = LearningWorkflow()
        self._learning_thread: Optional[threading.Thread] = None
        self._learning_args: Optional[tuple] = None

        p = self._communication_protocol
        partial_model = PartialModelCommand(self.state, self.stop, self.aggregator, p, self.learner)
        self.state.absorb_early_models = partial_model.absorb_early
        full_model = FullModelCommand(self.state, self.stop, self.aggregator, self.learner, p)
        self.state.adopt_early_full_model = full_model.adopt_early
        p.add_command(
            [
                StartLearningCommand(self.__start_learning_thread),
                StopLearningCommand(self.state, self.aggregator, self.learner),
                ModelInitializedCommand(self.state),
                VoteTrainSetCommand(self.state),
                ModelsAggregatedCommand(self.state),
                ModelsReadyCommand(self.state),
                MetricsCommand(self.state),
                InitModelCommand(self.state, self.stop, self.aggregator, self.learner, p),
                partial_model,
                full_model,
            ]
        )
        # wake waiting stages when the neighbour table or the aggregator changes
        hook = getattr(p, "on_neighbors_changed", None)
        if hook is not None:
            hook(self._on_neighbors_changed)
        self.aggregator.on_model_added(self.state.notify)

    # ------------------------------------------------------------------ neighbours
    def connect(self, addr: str) -> bool:
        self.assert_running(True)
        return self._communication_protocol.connect(addr)

    def get_neighbors(self, only_direct: bool = False) -> Dict[str, Any]:
        return self._communication_protocol.get_neighbors(only_direct)

    def disconnect(self, addr: str) -> None:
        self.assert_running(True)
        logger.info(self.addr, f"Removing {addr}...")
        self._communication_protocol.disconnect(addr, disconnect_msg=True)

    def _on_neighbors_changed(self) -> None:
        st = self.state
        if st.round is not None and st.train_set:
            # failure detector -> aggregation: a train-set member that is no longer a neighbour will not deliver its model
            try:
                alive = set(self._communication_protocol.get_neighbors(only_direct=False)) | {self.addr}
                for dead in [n for n in st.train_set if n not in alive]:
                    if self.aggregator.drop_node(dead):
                        st.train_set = [n for n in st.train_set if n != dead]
                        logger.info(self.addr, f"☠️ {dead} left during the round: aggregating without it.")
            except Exception:  # noqa: BLE001
                pass
        self.state.notify()
        if Settings.ALLOW_REJOIN and self.state.round is not None and self._learning_args is not None:
            self._invite_newcomers()

    def _invite_newcomers(self) -> None:
        """Rejoin: a direct neighbour that appears while we are learning and has announced nothing gets the
        experiment parameters (remaining rounds are implied by the current round); it will receive the model
        through the normal ``add_model`` diffusion of the round it enters in."""
        st = self.state
        try:
            direct = list(self._communication_protocol.get_neighbors(only_direct=True))
        except Exception:  # noqa: BLE001
            return
        rounds, epochs = self._learning_args  # type: ignore[misc]
        for n in direct:
            if n not in st.nei_status and n not in self._invited:
                self._invited.add(n)
                rnd = st.round
                if rnd is None:
                    return
                msg = self._communication_protocol.build_msg(StartLearningCommand.get_name(), [str(rounds), str(epochs), str(rnd)])
                msg["ttl"] = 0 if isinstance(msg, dict) else None
                with contextlib.suppress(Exception):
                    self._communication_protocol.send(n, msg, remove_on_error=False)

    # ------------------------------------------------------------------ lifecycle
    def assert_running(self, running: bool) -> None:
        if self.__running != running:
            raise NodeRunningException(f"Node is {'not ' if running else ''}running.")

    def start(self, wait: bool = False) -> None:
        self.assert_running(False)
        self.__running = True
        logger.register_node(self.addr, self.simulation)
        self._communication_protocol.start()
        if wait:
            self._communication_protocol.wait_for_termination()
            logger.info(self.addr, "Protocol terminated.")

    def stop(self) -> None:
        logger.info(self.addr, "Stopping node...")
        try:
            self.learner.interrupt_fit()
            self._communication_protocol.stop()
        except Exception:  # noqa: BLE001
            pass
        self.__running = False
        self.state.clear()
        with contextlib.suppress(Exception):
            logger.unregister_node(self.addr)

    # ------------------------------------------------------------------ learner plumbing
    def set_learner(self, learner: Learner) -> None:
        if self.state.round is not None:
            raise LearnerRunningException("Learner cannot be set after learning is started.")
        self.learner = learner

    def set_model(self, model: P2PFLModel) -> None:
        if self.state.round is not None:
            raise LearnerRunningException("Model cannot be set after learning is started.")
        self.learner.set_model(model)

    def set_data(self, data) -> None:
        if self.state.round is not None:
            raise LearnerRunningException("Data cannot be set after learning is started.")
        self.learner.set_data(data)

    def get_model(self) -> P2PFLModel:
        return self.learner.get_model()

    def get_data(self):
        return self.learner.get_data()

    # ------------------------------------------------------------------ learning control
    def set_start_learning(self, rounds: int = 1, epochs: int = 1) -> None:
        self.assert_running(True)
        if rounds < 1:
            raise ZeroRoundsException("Rounds must be greater than 0.")
        if self.state.round is not None:
            logger.info(self.addr, "Learning already started")
            return
        logger.info(self.addr, "🚀 Broadcasting start learning...")
        p = self._communication_protocol
        p.broadcast(p.build_msg(StartLearningCommand.get_name(), [str(rounds), str(epochs)]))
        with contextlib.suppress(RuntimeError):
            self.state.model_initialized_lock.release()  # the initiator owns the initial weights
        p.broadcast(p.build_msg(ModelInitializedCommand.get_name()))
        self.__start_learning_thread(rounds, epochs)

    def set_stop_learning(self) -> None:
        if self.state.round is not None:
            p = self._communication_protocol
            p.broadcast(p.build_msg(StopLearningCommand.get_name()))
            self.__stop_learning()
        else:
            logger.info(self.addr, "Learning already stopped")

    def __start_learning_thread(self, rounds: int, epochs: int, start_round: int = 0) -> None:
        with self.state.start_thread_lock:
            old = self._learning_thread
            winding_down = None
            if old is not None and old.is_alive():
                if not (getattr(old, "b2_experiment_set", False) and self.state.round is None):
                    return  # duplicated start (flood + direct invite) of the experiment that thread is running
                # that thread's experiment was stopped (its state is cleared) and it has not noticed yet - it is inside a
                # wait or a fit.  Dropping this start would leave the node out of the new experiment while everybody waits
                # VOTE_TIMEOUT for its ballot: the new learning thread starts as soon as the old one is gone.
                winding_down = old
            self._learning_args = (rounds, epochs)
            self._invited: set = set()
            t = threading.Thread(target=self.__start_learning, args=(rounds, epochs, start_round, winding_down), name="learning_thread-" + self.addr, daemon=True)
            self._learning_thread = t
        t.start()

    def __start_learning(self, rounds: int, epochs: int, start_round: int = 0, after: Optional[threading.Thread] = None) -> None:
        try:
            if after is not None:
                after.join(timeout=30)
            if start_round > 0:
                threading.current_thread().b2_experiment_set = True  # type: ignore[attr-defined]
                # rejoin: we enter a running experiment; the model comes with this round's add_model diffusion,
                # so we behave like a node outside the train set for our first round
                self.state.set_experiment(Settings.EXPERIMENT_NAME, rounds, round=start_round)
                self.learner.set_epochs(epochs)
                logger.experiment_started(self.addr, self.state.experiment)
                with contextlib.suppress(RuntimeError):
                    self.state.model_initialized_lock.release()
                self.state.rejoining = True
                # we do not vote in the round we enter: tell everybody, so nobody waits VOTE_TIMEOUT for our ballot
                p = self._communication_protocol
                with contextlib.suppress(Exception):
                    p.broadcast(p.build_msg(VoteTrainSetCommand.get_name(), [], round=start_round))
                first = "WaitAggregatedModelsStage"
                kw: Dict[str, Any] = {"first_stage": first}
            else:
                kw = {}
            self.learning_workflow.run(
                rounds=rounds, epochs=epochs, state=self.state, learner=self.learner, communication_protocol=self._communication_protocol, aggregator=self.aggregator, **kw
            )
        except Exception as e:  # noqa: BLE001
            if not self.__running:  # the node was stopped (or crashed on purpose) under its own learning thread
                logger.debug(self.addr, f"learning thread ended after stop: {type(e).__name__}: {e}")
                self.learning_workflow.finished = True
                return
            if self.state.round is None and getattr(threading.current_thread(), "b2_experiment_set", False):
                # the experiment was stopped under this thread (set_stop_learning clears the state at once; the thread was between
                # an early-stop check and its next use of the round counter): that is an early stop, not a reason to stop the NODE
                logger.debug(self.addr, f"learning thread ended by a stop: {type(e).__name__}: {e}")
                self.learning_workflow.finished = True
                return
            logger.error(self.addr, f"Error {type(e).__name__}: {e}\n{traceback.format_exc()}")
            self.stop()

    def __stop_learning(self) -> None:
        logger.info(self.addr, "Stopping learning")
        self.learner.interrupt_fit()
        self.aggregator.clear()
        self.state.clear()
        logger.experiment_finished(self.addr)
        with contextlib.suppress(Exception):
            self.state.wait_votes_ready_lock.release()
