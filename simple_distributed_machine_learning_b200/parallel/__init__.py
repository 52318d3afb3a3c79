"""Pipeline / data / tensor parallel runtime: topology, schedules, executors, transports, fused
collectives.  Replaces the reference's RPC + distributed-autograd coupling of its two stages
(/root/reference/simple_distributed.py:33-50, :100-113)."""
