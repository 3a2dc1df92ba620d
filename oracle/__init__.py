"""CPU oracle of the pyhgf belief-propagation path -- TEST INFRASTRUCTURE ONLY.

See oracle/README.md.  Import from tests/, __graft_entry__.smoke() and the CPU
baseline legs of bench.py only; the product package never imports this.
"""
