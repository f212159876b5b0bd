"""Command line tool: ``python -m myfyp_b200 experiment list|run|help`` (reference: p2pfl/cli.py:137-238).

``login`` / ``remote`` / ``launch`` answer "not implemented yet" exactly like the reference (cli.py:71-95).  Experiments are
the scripts in ``myfyp_b200/examples``; ``run`` executes one as a subprocess and forwards its flags.  ``build`` and
``bench`` are additions: compile the native library, run the headline benchmark.
"""

from __future__ import annotations

import os
import subprocess
import sys
from typing import Dict, List, Optional

import typer
from rich.console import Console
from rich.panel import Panel
from rich.table import Table

console = Console()
app = typer.Typer(rich_markup_mode="rich", help="[bold blue]myfyp_b200[/bold blue] | Blackwell-native peer-to-peer federated learning")
exp_app = typer.Typer(help="Run example experiments.")
app.add_typer(exp_app, name="experiment")

EXAMPLES_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "examples")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _examples() -> Dict[str, str]:
    out = {}
    for fn in sorted(os.listdir(EXAMPLES_DIR)):
        if fn.endswith(".py") and not fn.startswith("__"):
            with open(os.path.join(EXAMPLES_DIR, fn)) as f:
                parts = f.read().split('"""')
            out[fn[:-3]] = parts[1].strip() if len(parts) > 1 else ""
    return out


def _not_yet() -> None:
    console.print(":sweat_smile: [bold yellow]Not implemented yet![/bold yellow]")


@app.command()
def login(token: str = typer.Option("", help="API token")) -> None:
    """Authenticate with the web platform."""
    _not_yet()


@app.command()
def remote() -> None:
    """Interact with a remote node."""
    _not_yet()


@app.command()
def launch(
    address: str = typer.Option("127.0.0.1", help="ip[:port], [v6]:port, unix:///path (gRPC) or uds:///path.sock"),
    connect: Optional[List[str]] = typer.Option(None, help="address of a node to connect to (repeatable)"),
    model: str = typer.Option("mlp", help="mlp | cnn | resnet18"),
    aggregator: str = typer.Option("fedavg", help="fedavg | fedmedian | krum | scaffold | fedprox"),
    samples: int = typer.Option(600, help="synthetic training samples of this node (real data: use the Python API)"),
    seed: int = 0,
    rounds: int = typer.Option(0, help="> 0: start learning (this node initiates) once connected, and exit when it finishes"),
    epochs: int = 1,
    ssl: bool = False,
    settings: Optional[str] = typer.Option(None, help="JSON / YAML file of Settings values"),
) -> None:
    """Launch one node of a federation in this process and keep it running (the reference leaves this command unimplemented,
    cli.py:87-95).  Without --rounds it serves until interrupted; other nodes connect to it and one of them starts the learning."""
    import time

    from myfyp_b200.learning.aggregators import FedAvg, FedMedian, FedProx, Krum, Scaffold
    from myfyp_b200.learning.dataset import P2PFLDataset
    from myfyp_b200.models import cnn_model, mlp_model, resnet18_model
    from myfyp_b200.node import Node
    from myfyp_b200.settings import Settings
    from myfyp_b200.utils.utils import wait_to_finish

    Settings.update_from_env()
    if settings:
        Settings.load_file(settings)
    Settings.USE_SSL = ssl
    mk = {"mlp": mlp_model, "cnn": cnn_model, "resnet18": resnet18_model}[model]
    agg = {"fedavg": FedAvg, "fedmedian": FedMedian, "krum": lambda: Krum(f=1), "scaffold": lambda: Scaffold(global_lr=1.0), "fedprox": lambda: FedProx(mu=0.01)}[aggregator]()
    shape = (28, 28) if model == "mlp" else (3, 32, 32)
    data = P2PFLDataset.synthetic(shape, 10, samples, max(64, samples // 5), seed=seed)
    protocol = None
    if address.startswith("uds://"):
        from myfyp_b200.communication.protocols.uds.uds_communication_protocol import UdsCommunicationProtocol as protocol  # noqa: N813
    node = Node(mk(seed=seed), data, address=address, aggregator=agg, protocol=protocol)
    node.start()
    console.print(f":satellite: node up at [green]{node.addr}[/green]")
    try:
        for peer in connect or []:
            deadline = time.time() + 60
            while not node.connect(peer):
                if time.time() > deadline:
                    raise typer.Exit(code=1)
                time.sleep(0.5)
            console.print(f":link: connected to {peer}")
        if rounds > 0:
            time.sleep(2 * Settings.HEARTBEAT_PERIOD)  # let the heartbeats flood before candidates are counted
            node.set_start_learning(rounds=rounds, epochs=epochs)
            wait_to_finish([node], timeout=24 * 3600)
            console.print(":checkered_flag: learning finished")
        else:
            while True:
                time.sleep(1.0)
    except KeyboardInterrupt:
        pass
    finally:
        node.stop()


@app.command()
def build(force: bool = False, verbose: bool = False) -> None:
    """Compile the sm_100a kernels into myfyp_b200/lib/libmyfyp_b200.so (needs nvcc, not a GPU)."""
    from myfyp_b200.build import build as _build

    console.print(f":hammer: built [green]{_build(force, verbose)}[/green]")


@app.command(context_settings={"allow_extra_args": True, "ignore_unknown_options": True})
def bench(ctx: typer.Context) -> None:
    """Run bench.py (FL rounds/sec); extra flags are forwarded, e.g. --steps 10 --warmup 3."""
    raise typer.Exit(subprocess.call([sys.executable, os.path.join(ROOT, "bench.py"), *ctx.args]))


@exp_app.command("list")
def list_examples() -> None:
    """List the available examples."""
    table = Table("Name", "Description", title="Examples", show_lines=True)
    for name, doc in _examples().items():
        table.add_row(name, doc.split("\n\n")[0])
    console.print(table)


@exp_app.command("help")
def help_example(example: str) -> None:
    """Show the flags of an example."""
    _check(example)
    raise typer.Exit(subprocess.call([sys.executable, os.path.join(EXAMPLES_DIR, example + ".py"), "--help"]))


@exp_app.command("run", context_settings={"allow_extra_args": True, "ignore_unknown_options": True})
def run_example(ctx: typer.Context, example: str) -> None:
    """Run an example; unknown options are passed through to it."""
    _check(example)
    console.print(Panel(f":rocket: running [bold]{example}[/bold] {' '.join(ctx.args)}"))
    raise typer.Exit(subprocess.call([sys.executable, os.path.join(EXAMPLES_DIR, example + ".py"), *ctx.args]))


def _check(example: str) -> None:
    if example not in _examples():
        console.print(Panel(f":x: example [yellow]{example}[/yellow] not found; see [green]experiment list[/green]", title="[red]Example not found[/red]"))
        raise typer.Exit(code=1)


def main(argv: Optional[List[str]] = None) -> None:
    app(args=argv)


if __name__ == "__main__":
    main()
