"""CPU oracle for the gnnepcsaft message-passing hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``gnnepcsaft_b200/`` imports this
package; only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may.

Parity status: **parity unpinned** -- the reference ships no tests, golden
vectors, checkpoints or processed datasets for this path, and the arithmetic
lives in two un-vendored, un-pinned third-party packages (``torch_geometric``
and ``ogb``) that cannot be imported in the build container or on the GPU box.
``pyg_restated`` is a restatement of their published pure-torch semantics,
anchored on the reference's own call sites (``gnnepcsaft/train/models.py``).
"""
