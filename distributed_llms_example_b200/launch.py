"""Runs an unmodified reference script with the B200 hot path attached.

    torchrun --nproc-per-node=N -m distributed_llms_example_b200.launch train-torchrun.py {parameters}
    python -m accelerate.commands.launch ... -m distributed_llms_example_b200.launch train-accelerator.py ...
    python -m distributed_llms_example_b200.launch train-task.py {parameters}

(the launch lines of valohai.yaml:7,33,59 with ``-m distributed_llms_example_b200.launch`` put in front of
the script name). It applies patch.apply() — CrossEntropyLoss, torch.optim.AdamW and the DDP comm
hook — and, because ``average_gradients`` (train-task.py:65-69) is defined by the script itself, it
rebinds that name in the script's namespace immediately after the script's own ``def``: the file on
disk is not edited, its training logic runs as written.

Limitation stated in README/DESIGN: on the build image the scripts stop at ``import valohai`` /
``accelerate`` / ``evaluate`` (not installable offline); this launcher does not fake those packages.
"""
from __future__ import annotations

import ast
import os
import sys
from typing import Callable, Mapping, Optional


def _average_gradients(model) -> None:
    """Late-bound so that importing the launcher does not need CUDA."""
    from .comm import average_gradients

    average_gradients(model)


DEFAULT_REBINDS: dict[str, Callable] = {"average_gradients": _average_gradients}


def rebind_after_definition(source: str, filename: str, names: Mapping[str, Callable]) -> ast.Module:
    """Parse ``source`` and, after every top-level ``def <name>`` with <name> in ``names``, insert
    ``<name> = __b200_rebinds__["<name>"]``."""
    tree = ast.parse(source, filename=filename)
    body: list[ast.stmt] = []
    for node in tree.body:
        body.append(node)
        if isinstance(node, (ast.FunctionDef, ast.AsyncFunctionDef)) and node.name in names:
            assign = ast.parse(f'{node.name} = __b200_rebinds__["{node.name}"]').body[0]
            ast.copy_location(assign, node)
            body.append(assign)
    tree.body = body
    return ast.fix_missing_locations(tree)


def run_script(path: str, argv: Optional[list[str]] = None, rebinds: Optional[Mapping[str, Callable]] = None,
               apply_patches: bool = True, run_name: str = "__main__") -> dict:
    """Execute the script at ``path`` as ``__main__`` with the plug-ins attached; returns its globals."""
    rebinds = DEFAULT_REBINDS if rebinds is None else rebinds
    if apply_patches:
        from . import patch

        patch.apply()
    path = os.path.abspath(path)
    with open(path, "r", encoding="utf-8") as f:
        source = f.read()
    code = compile(rebind_after_definition(source, path, rebinds), path, "exec")
    globs = {"__name__": run_name, "__file__": path, "__b200_rebinds__": dict(rebinds),
             "__builtins__": __builtins__}
    old_argv, old_path = sys.argv, list(sys.path)
    sys.argv = [path] + list(argv or [])
    sys.path.insert(0, os.path.dirname(path))  # the scripts do ``import helpers``
    try:
        exec(code, globs)
    finally:
        sys.argv, sys.path[:] = old_argv, old_path
    return globs


def main() -> None:
    if len(sys.argv) < 2:
        raise SystemExit("usage: python -m distributed_llms_example_b200.launch <script.py> [script args...]")
    run_script(sys.argv[1], sys.argv[2:])


if __name__ == "__main__":
    main()
