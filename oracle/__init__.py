"""CPU oracle for the UQ inference hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this package; the product (the ``uncertaintyquantification-sleepapnea-1dcnn_b200``
package) never does and fails loudly when its CUDA library is missing.

Parity status
-------------
* Forward passes (``network.py``): **parity unpinned**.  The arithmetic lives in TensorFlow/Keras 2.12.0
  (reference README.md:79), which is neither vendored in /root/reference nor installable in this image,
  and the reference has no tests or golden vectors.  ``network.py`` restates the
  published Keras layer semantics and anchors on the reference's own call sites
  (uq_techniques.py:22, :29; cnn_baseline_train.py:55-94).
* Reductions (``reductions.py``, ``patient.py``): pinned.  They are checked against outputs of the
  reference's *own* NumPy/SciPy/pandas code, imported from /root/reference in the build container with
  ``tensorflow``/``matplotlib`` stubbed (``reference_loader.py``); the outputs are committed under
  ``tests/golden/`` together with the generating script ``tests/golden/make_golden.py``.
"""
