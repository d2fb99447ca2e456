"""TEST INFRASTRUCTURE ONLY: CPU restatements of the reference algorithms (the parity oracle).

Importers allowed: tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference legs.
"""
