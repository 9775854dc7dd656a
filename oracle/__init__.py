"""CPU oracle — TEST INFRASTRUCTURE ONLY.

Restates the reference path (services/supervisor.go:137-374 of
SneaksAndData/nexus-supervisor) on the CPU.  Import it only from tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs;
the product package never does.
"""
