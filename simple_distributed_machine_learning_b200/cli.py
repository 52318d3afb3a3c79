"""CLI glue: parse flags (reference contract, ``utils/config.py``) and run this rank."""
from __future__ import annotations

from typing import Optional, Sequence

from .utils.config import parse_args


def main(argv: Optional[Sequence[str]] = None) -> None:
    cfg = parse_args(argv)
    from .train import run
    run(cfg)


if __name__ == "__main__":
    main()
