"""CPU oracle for the CTC hot path -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import this package.  The product package (``ctc-attention-mispronunciation_b200``) never does; it fails loudly
when its CUDA library is missing instead of falling back to anything here.

``oracle.api`` wraps ``mdd_oracle.c`` (a C restatement of the reference algorithms, each function citing the
reference file:line it follows) through ctypes.  ``oracle.ref_loader`` imports the real reference from
``/root/reference`` (present only in the build container) and ``oracle.gen_golden`` uses it to write the
fixtures under ``tests/golden/`` that pin the restatement.
"""
