"""CPU oracle: TEST INFRASTRUCTURE ONLY. Importable from tests/, __graft_entry__.smoke() and bench.py's CPU legs; never from the product package."""
