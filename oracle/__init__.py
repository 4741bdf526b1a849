"""CPU oracle for the stitch hot path (TEST INFRASTRUCTURE ONLY).

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import or execute it, and only as the checker / the timed
CPU baseline -- never on the product path.

Parity status
-------------
* layout  : PINNED BY THE REFERENCE'S EXECUTED SOURCE.  ``layout_ref.py`` is a
            restatement of the reference's double arithmetic
            (miniprogram-stitch/miniprogram/pages/index/index.js:1211-1213,
            1243-1386, 1432-1433, 1522-1557; utils/canvas.js:153-202).  The
            reference ships no tests, golden vectors or fixtures and no JS engine
            exists in this image, so ``oracle/jsmini.py`` -- a JavaScript-subset
            interpreter -- runs those lines VERBATIM (sliced by line number from
            /root/reference) under a ~20-line page / wx / canvas stub;
            ``gen_golden.py`` writes tests/golden/layout_golden.json from those
            runs and tests/test_jsmini.py holds the restatement, the C++ twin
            and the orientation table against them (33 goldens + a 250-case
            hypothesis sweep, bit-exact doubles).
* pixels  : "parity unpinned".  The reference delegates resampling to the closed
            WeChat Canvas2D runtime (ctx.drawImage, utils/canvas.js:156-199);
            the filter used here ("AB1") is a stated
            definition, cross-checked against cv2.INTER_AREA / INTER_LINEAR.
"""
