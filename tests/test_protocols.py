"""Protocol tests parametrised over both transports (reference: test/communication/communication_test.py)."""

import time

import pytest

from myfyp_b200.communication.commands.command import Command
from myfyp_b200.communication.protocols.exceptions import CommunicationError, NeighborNotConnectedError, ProtocolNotStartedError
from myfyp_b200.communication.protocols.grpc.grpc_communication_protocol import GrpcCommunicationProtocol
from myfyp_b200.communication.protocols.memory.memory_communication_protocol import InMemoryCommunicationProtocol
from myfyp_b200.communication.protocols.uds.uds_communication_protocol import UdsCommunicationProtocol
from myfyp_b200.settings import Settings
from myfyp_b200.utils.utils import set_test_settings


class GrpcOverSockets(GrpcCommunicationProtocol):
    """gRPC with the in-process shortcut off: every message really crosses a socket and the hand-written codec."""

    def __init__(self, addr="127.0.0.1", commands=None):
        Settings.GRPC_LOCAL_SHORTCUT = False
        super().__init__(addr, commands)


class UdsOverSockets(UdsCommunicationProtocol):
    """Unix-socket protocol with the in-process shortcut off (frames really cross the socket)."""

    def __init__(self, addr=None, commands=None):
        Settings.GRPC_LOCAL_SHORTCUT = False
        super().__init__(addr, commands)


PROTOCOLS = [GrpcCommunicationProtocol, GrpcOverSockets, InMemoryCommunicationProtocol, UdsOverSockets]


class MockCommand(Command):
    def __init__(self):
        self.calls = []

    @staticmethod
    def get_name():
        return "mock"

    def execute(self, source, round, *args, **kwargs):
        self.calls.append((source, round, args, kwargs))


@pytest.fixture(autouse=True)
def _settings():
    set_test_settings()
    from myfyp_b200.management.logger import logger

    logger.set_level("ERROR")
    yield


@pytest.mark.parametrize("proto", PROTOCOLS)
def test_connect_to_invalid_address(proto):
    p = proto()
    p.start()
    try:
        assert p.connect("google.es:80") is False
        assert p.connect("holadani.holaenrique") is False
        assert p.connect(p.get_address()) is False  # itself
        assert len(p.get_neighbors()) == 0
    finally:
        p.stop()


@pytest.mark.parametrize("proto", PROTOCOLS)
def test_errors_and_command_dispatch(proto):
    cmd1, cmd2 = MockCommand(), MockCommand()
    p1, p2 = proto(commands=[cmd1]), proto()
    p2.add_command(cmd2)
    with pytest.raises(ProtocolNotStartedError):
        p1.connect("x")
    with pytest.raises(ProtocolNotStartedError):
        p1.get_neighbors()
    p1.start()
    p2.start()
    try:
        with pytest.raises(NeighborNotConnectedError):
            p1.send(p2.get_address(), p1.build_msg("mock"), raise_error=True)
        assert p1.connect(p2.get_address())
        assert p2.get_address() in p1.get_neighbors(only_direct=True) and p1.get_address() in p2.get_neighbors(only_direct=True)
        with pytest.raises(CommunicationError):
            p1.send(p2.get_address(), p1.build_msg("no_such_command"), raise_error=True, remove_on_error=False)
        p1.send(p2.get_address(), p1.build_msg("mock", ["a", "b"], round=3), raise_error=True)
        assert cmd2.calls[-1][:3] == (p1.get_address(), 3, ("a", "b"))
        p1.send(p2.get_address(), p1.build_weights("mock", 7, b"\x00payload\xff", ["c1", "c2"], 42), raise_error=True)
        src, rnd, args, kw = cmd2.calls[-1]
        assert (src, rnd, args) == (p1.get_address(), 7, ()) and bytes(kw["weights"]) == b"\x00payload\xff"
        assert list(kw["contributors"]) == ["c1", "c2"] and kw["num_samples"] == 42
        p2.broadcast(p2.build_msg("mock", ["x"]))
        deadline = time.time() + 2  # a broadcast does not wait for the peers' verdicts (one-way frames on the socket protocol)
        while time.time() < deadline and not (cmd1.calls and cmd1.calls[-1][2] == ("x",)):
            time.sleep(0.005)
        assert cmd1.calls[-1][2] == ("x",)
    finally:
        p1.stop()
        p2.stop()


@pytest.mark.parametrize("proto", PROTOCOLS)
def test_neighbor_discovery_by_heartbeat_flood_and_disconnect(proto):
    n = 5
    ps = [proto() for _ in range(n)]
    for p in ps:
        p.start()
    try:
        for i in range(n - 1):  # a line: only heartbeats can teach the ends about each other
            assert ps[i].connect(ps[i + 1].get_address())
        deadline = time.time() + 8
        while time.time() < deadline and not all(len(p.get_neighbors(only_direct=False)) == n - 1 for p in ps):
            time.sleep(0.05)
        assert all(len(p.get_neighbors(only_direct=False)) == n - 1 for p in ps)
        assert len(ps[0].get_neighbors(only_direct=True)) == 1 and len(ps[2].get_neighbors(only_direct=True)) == 2
        ps[0].disconnect(ps[1].get_address())
        assert ps[1].get_address() not in ps[0].get_neighbors(only_direct=True)
        deadline = time.time() + 3
        while time.time() < deadline and ps[0].get_address() in ps[1].get_neighbors(only_direct=True):
            time.sleep(0.05)
        assert ps[0].get_address() not in ps[1].get_neighbors(only_direct=True)
    finally:
        for p in ps:
            p.stop()


@pytest.mark.parametrize("proto", PROTOCOLS)
def test_abrupt_stop_is_detected(proto):
    """A peer that dies without saying goodbye is evicted within 5 s (heartbeat 0.5 s / timeout 2 s)."""
    p1, p2 = proto(), proto()
    p1.start()
    p2.start()
    try:
        assert p1.connect(p2.get_address())
        time.sleep(0.3)
        # kill p2 without the polite disconnect
        p2._heartbeater.stop()
        p2._gossiper.stop()
        p2._server.stop()
        t0 = time.time()
        while time.time() - t0 < 5 and len(p1.get_neighbors()) > 0:
            time.sleep(0.05)
        assert len(p1.get_neighbors()) == 0
    finally:
        p1.stop()


@pytest.mark.parametrize("proto", PROTOCOLS)
@pytest.mark.parametrize("goodbye", [True, False])
def test_a_death_reaches_nodes_that_know_the_dead_one_only_through_the_flood(proto, goodbye):
    """Line a - b - c: ``a`` knows ``c`` only from c's flooded beats.  When c goes (an orderly stop floods its own ``node_down``;
    after a crash its direct neighbour b notices the failed beat and floods one), ``a`` drops it within a couple of beat periods
    instead of waiting out HEARTBEAT_TIMEOUT (set to 30 s here) - a vote needs every known node's ballot."""
    from myfyp_b200.settings import Settings

    old = Settings.HEARTBEAT_PERIOD, Settings.HEARTBEAT_TIMEOUT
    Settings.HEARTBEAT_PERIOD, Settings.HEARTBEAT_TIMEOUT = 0.2, 30.0
    a, b, c = proto(), proto(), proto()
    try:
        for p in (a, b, c):
            p.start()
        assert a.connect(b.get_address()) and b.connect(c.get_address())
        t0 = time.time()
        while time.time() - t0 < 5 and c.get_address() not in a.get_neighbors(only_direct=False):
            time.sleep(0.02)
        assert c.get_address() in a.get_neighbors(only_direct=False) and c.get_address() not in a.get_neighbors(only_direct=True)
        if goodbye:
            c.stop()
        else:  # crash: no disconnect, no goodbye
            c._heartbeater.stop()
            c._gossiper.stop()
            c._server.stop()
        t0 = time.time()
        while time.time() - t0 < 4 and c.get_address() in a.get_neighbors(only_direct=False):
            time.sleep(0.02)
        assert c.get_address() not in a.get_neighbors(only_direct=False), "a still waits for c's beats to time out"
        assert b.get_address() in a.get_neighbors(only_direct=True)
    finally:
        Settings.HEARTBEAT_PERIOD, Settings.HEARTBEAT_TIMEOUT = old
        for p in (a, b):
            p.stop()
        if not goodbye:
            try:
                c._neighbors.clear_neighbors()
            except Exception:  # noqa: BLE001
                pass


def test_a_peer_that_answers_with_an_error_is_not_declared_dead():
    """``node_down`` is for peers that are GONE (no server, refused / broken / silent connection).  A peer whose handler raised has
    answered: its direct neighbour may drop the link (the reference's behaviour), but nobody else is told to forget the node."""
    from myfyp_b200.communication.protocols.exceptions import CommunicationError, NeighborNotConnectedError, peer_looks_dead

    assert peer_looks_dead(ConnectionRefusedError()) and peer_looks_dead(TimeoutError()) and peer_looks_dead(CommunicationError("Error while sending a message: 'x': 'server not running'"))
    assert not peer_looks_dead(CommunicationError("Error while sending a message: 'mock': 'Error while processing command: mock []: boom'"))
    assert not peer_looks_dead(NeighborNotConnectedError("no stub")) and not peer_looks_dead(ValueError("x"))

    class Boom(MockCommand):
        def execute(self, source, round, *args, **kwargs):
            raise RuntimeError("boom")

    a, b, c = (InMemoryCommunicationProtocol() for _ in range(3))
    try:
        for p in (a, b, c):
            p.start()
        assert a.connect(b.get_address()) and b.connect(c.get_address())
        t0 = time.time()
        while time.time() - t0 < 5 and c.get_address() not in a.get_neighbors(only_direct=False):
            time.sleep(0.02)
        c.add_command(Boom())
        b.send(c.get_address(), b.build_msg("mock"))  # c answers with an error: b drops its link to c ...
        assert c.get_address() not in b.get_neighbors(only_direct=True)
        time.sleep(0.3)
        assert c.get_address() in a.get_neighbors(only_direct=False)  # ... but a was not told that c is dead
    finally:
        for p in (a, b, c):
            p.stop()


def test_stopping_one_memory_node_leaves_the_others_registered():
    a, b, c = (InMemoryCommunicationProtocol() for _ in range(3))
    for p in (a, b, c):
        p.start()
    a.connect(b.get_address())
    c.stop()  # the reference wipes the whole registry here
    assert a.connect(c.get_address()) is False
    cmd = MockCommand()
    b.add_command(cmd)
    a.send(b.get_address(), a.build_msg("mock"), raise_error=True)
    assert len(cmd.calls) == 1
    a.stop()
    b.stop()


def test_flood_dedup_and_ttl():
    """A control message reaches every node of a ring exactly once."""
    n = 6
    cmds = [MockCommand() for _ in range(n)]
    ps = [InMemoryCommunicationProtocol(commands=[cmds[i]]) for i in range(n)]
    for p in ps:
        p.start()
    try:
        for i in range(n):
            ps[i].connect(ps[(i + 1) % n].get_address())
        ps[0].broadcast(ps[0].build_msg("mock", ["hello"]))
        deadline = time.time() + 3
        while time.time() < deadline and not all(len(c.calls) == 1 for c in cmds[1:]):
            time.sleep(0.02)
        time.sleep(0.2)
        assert [len(c.calls) for c in cmds] == [0] + [1] * (n - 1)
    finally:
        for p in ps:
            p.stop()


def test_wire_codec_matches_protobuf():
    """The hand-written codec produces/consumes exactly what google.protobuf does for node.proto."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

    from myfyp_b200.communication.protocols.grpc import wire

    fd = descriptor_pb2.FileDescriptorProto(name="node_test.proto", package="nodetest", syntax="proto3")
    T = descriptor_pb2.FieldDescriptorProto

    def add(msg, name, num, typ, label=T.LABEL_OPTIONAL, type_name=None, oneof=None, optional=False):
        f = msg.field.add(name=name, number=num, type=typ, label=label)
        if type_name:
            f.type_name = type_name
        if oneof is not None:
            f.oneof_index = oneof
        if optional:
            f.proto3_optional = True

    m = fd.message_type.add(name="Message")
    add(m, "ttl", 1, T.TYPE_INT32)
    add(m, "hash", 2, T.TYPE_INT64)
    add(m, "args", 3, T.TYPE_STRING, T.LABEL_REPEATED)
    w = fd.message_type.add(name="Weights")
    add(w, "weights", 1, T.TYPE_BYTES)
    add(w, "contributors", 2, T.TYPE_STRING, T.LABEL_REPEATED)
    add(w, "num_samples", 3, T.TYPE_INT32)
    r = fd.message_type.add(name="RootMessage")
    r.oneof_decl.add(name="payload_type")
    r.oneof_decl.add(name="_round")
    add(r, "source", 1, T.TYPE_STRING)
    add(r, "round", 2, T.TYPE_INT32, oneof=1, optional=True)
    add(r, "cmd", 3, T.TYPE_STRING)
    add(r, "message", 4, T.TYPE_MESSAGE, type_name=".nodetest.Message", oneof=0)
    add(r, "weights", 5, T.TYPE_MESSAGE, type_name=".nodetest.Weights", oneof=0)
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fd)
    Root = message_factory.GetMessageClass(pool.FindMessageTypeByName("nodetest.RootMessage"))

    msg = {"source": "127.0.0.1:5000", "round": -1, "cmd": "vote_train_set", "ttl": 10, "hash": -1234567890123456789, "args": ["a", "7", "ünï"]}
    pb = Root()
    pb.ParseFromString(wire.encode_root(msg))
    assert (pb.source, pb.round, pb.cmd, pb.message.ttl, pb.message.hash, list(pb.message.args)) == ("127.0.0.1:5000", -1, "vote_train_set", 10, -1234567890123456789, ["a", "7", "ünï"])
    back = wire.decode_root(pb.SerializeToString())
    assert {k: back[k] for k in msg} == msg

    wmsg = {"source": "s", "round": 4, "cmd": "partial_model", "weights": bytes(range(256)) * 5, "contributors": ["x", "y"], "weight": 123}
    pb = Root()
    pb.ParseFromString(wire.encode_root(wmsg))
    assert pb.WhichOneof("payload_type") == "weights" and pb.weights.weights == wmsg["weights"] and pb.weights.num_samples == 123
    back = wire.decode_root(pb.SerializeToString())
    assert back["weights"] == wmsg["weights"] and back["contributors"] == ["x", "y"] and back["weight"] == 123 and back["round"] == 4
    assert wire.decode_response(wire.encode_response("boom")) == "boom" and wire.decode_response(wire.encode_response(None)) == ""
    assert wire.decode_handshake(wire.encode_handshake("unix:///tmp/a.sock")) == "unix:///tmp/a.sock"


def test_address_parser():
    from myfyp_b200.communication.protocols.grpc.address import AddressParser

    assert AddressParser("127.0.0.1:6000").get_parsed_address() == "127.0.0.1:6000"
    assert AddressParser("[::1]:6000").get_parsed_address() == "[::1]:6000"
    a = AddressParser("127.0.0.1")
    assert a.host == "127.0.0.1" and 1 <= a.port <= 65535
    assert AddressParser("unix:///tmp/x.sock").get_parsed_address() == "unix:///tmp/x.sock"
    with pytest.raises(ValueError):
        AddressParser("not an address:99999999").get_parsed_address()


def test_grpc_mutual_tls(tmp_path):
    """gRPC over mTLS with certificates from certificates/gen-certs.sh (the reference's CI generates its own the same way)."""
    import os
    import shutil
    import subprocess

    if shutil.which("openssl") is None:
        pytest.skip("openssl not available")
    script = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "myfyp_b200", "certificates", "gen-certs.sh")
    shutil.copy(script, tmp_path / "gen-certs.sh")
    subprocess.run(["bash", str(tmp_path / "gen-certs.sh")], check=True, capture_output=True)
    Settings.USE_SSL = True
    Settings.GRPC_LOCAL_SHORTCUT = False
    Settings.CA_CRT, Settings.SERVER_CRT, Settings.SERVER_KEY = str(tmp_path / "ca.crt"), str(tmp_path / "server.crt"), str(tmp_path / "server.key")
    Settings.CLIENT_CRT, Settings.CLIENT_KEY = str(tmp_path / "client.crt"), str(tmp_path / "client.key")
    cmd = MockCommand()
    p1, p2 = GrpcCommunicationProtocol("127.0.0.1"), GrpcCommunicationProtocol("127.0.0.1", commands=[cmd])
    p1.start()
    p2.start()
    try:
        assert p1.connect(p2.get_address())
        p1.send(p2.get_address(), p1.build_msg("mock", ["secret"]), raise_error=True)
        assert cmd.calls[-1][2] == ("secret",)
        # a client without certificates cannot talk to the TLS server
        import grpc

        from myfyp_b200.communication.protocols.grpc import wire

        ch = grpc.insecure_channel(p2.get_address())
        call = ch.unary_unary("/node.NodeServices/handshake", request_serializer=wire.encode_handshake, response_deserializer=wire.decode_response)
        with pytest.raises(grpc.RpcError):
            call("intruder", timeout=1.0)
        ch.close()
    finally:
        p1.stop()
        p2.stop()
