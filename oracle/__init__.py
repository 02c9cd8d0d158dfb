"""CPU oracle for the motion-module + UnZipLoRA attention hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``video_style_transfer_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs do, and there only as the checker / the timed CPU baseline.

Parity status: the reference (tanmud/video_style_transfer) ships no tests and no golden
vectors.  Every restatement in ``oracle/hotpath.py`` that has an importable
reference twin is pinned against the reference's own modules executed in the build
container by ``oracle/make_golden.py`` (outputs committed under ``tests/golden/``).
The diffusers-owned motion block (GroupNorm5D -> proj_in -> BasicTransformerBlock ->
proj_out) is absent from the reference tree and from this image: that one restatement is
"parity unpinned" (recalled structure, built from the pinned pieces).
"""
