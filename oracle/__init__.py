"""CPU oracle for the stitch hot path (TEST INFRASTRUCTURE ONLY).

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import or execute it, and only as the checker / the timed
CPU baseline -- never on the product path.

Parity status
-------------
* layout  : restatement of the reference's double arithmetic
            (miniprogram-stitch/miniprogram/pages/index/index.js:1251-1386,
            1432-1433, 1522-1554).  The reference ships no tests, golden
            vectors or fixtures; see ``oracle/jsmini.py`` for how the
            restatement is pinned against the reference's own source text.
* pixels  : "parity unpinned".  The reference delegates resampling to the closed
            WeChat Canvas2D runtime (ctx.drawImage, utils/canvas.js:156-199);
            the filter used here ("AB1") is a stated
            definition, cross-checked against cv2.INTER_AREA / INTER_LINEAR.
"""
