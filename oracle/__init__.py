"""CPU oracle for the BERT-QA hot path.  TEST INFRASTRUCTURE ONLY.

Nothing in the product package imports this directory; only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / reference legs may use it (as the checker, never as the thing
measured or shipped).
"""
