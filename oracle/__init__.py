"""TEST INFRASTRUCTURE ONLY.

CPU restatement (PyTorch fp32 on the host, plus one plain-C file for DCN) of the
reference's stereo-matching hot path.  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import this
package.  The product package ``stereoformer_b200`` never does: it fails loudly
when its CUDA library is missing.
"""
