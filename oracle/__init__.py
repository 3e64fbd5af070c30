"""TEST INFRASTRUCTURE ONLY — CPU restatement ("oracle") of the reference's hot path.

Importable only from ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs.  The product package ``options_in_simulation_b200`` never imports it.
"""
