#!/usr/bin/env python
"""Command-line entry point, same invocation as the reference script:

    python simple_distributed.py --rank=0 --world_size=2 --interface=eth0 \
        --master_addr=192.168.10.128 --master_port=2308

(/root/reference/simple_distributed.py:138-186; its README names the file
``distributed.py``, the actual file is ``simple_distributed.py``.)  One process per
rank, started by hand or by any launcher; on a GPU box each rank drives one B200.
See ``--help`` for the added flags (--pp/--dp/--tp, --microbatches, --schedule, ...).
"""
from simple_distributed_machine_learning_b200.cli import main

if __name__ == "__main__":
    main()
