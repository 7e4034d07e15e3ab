"""CPU oracle of the hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package; it is the checker, never the product.  The product
(imaginaire_b200/) never imports it and fails loudly when its CUDA library is
missing.
"""
