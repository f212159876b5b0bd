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
def launch() -> None:
    """Launch a node."""
    _not_yet()


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
