"""oracle -- TEST INFRASTRUCTURE ONLY.

CPU checkers for the compressed-genotype hot path:

* ``oracle.cport``  : ctypes wrapper around ``liboracle.so`` (our plain-C restatement, oracle.c)
* ``oracle.ref``    : ctypes wrapper around ``_ref/ref_miraculix.so`` (the UNMODIFIED reference CPU
                      "5codes" library compiled by oracle/Makefile from /root/reference)
* ``oracle.simulate``: seeded numpy generator of simulated PLINK inputs

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
legs may import this package; the product (``miraculix_b200``) never does.
"""
