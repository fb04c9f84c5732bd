"""CPU oracle for the layer-range forward + sampling path (TEST INFRASTRUCTURE ONLY).

Nothing under ``xotorch_support_jetson_b200/`` may import this package.  Only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` use it, and only as the checker / the timed CPU baseline.
"""
