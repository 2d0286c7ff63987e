"""TEST INFRASTRUCTURE ONLY.

CPU restatement ("oracle") of the GNN_Hex hot path.  Nothing in ``gnn_hex_b200/`` may import
this package: only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs use it, and only as the checker / timed CPU baseline.
"""
