"""CPU oracle for the MuZero-Go self-play hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``muzero_go_b200/`` may import this
package.  Allowed importers: ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py``.

Contents
  gymgo.py      literal numpy/scipy restatement of huangeddie/GymGo
                (pinned commit 3a3d97b; source ABSENT from /root/reference,
                so rules parity is *unpinned by the reference*)
  gymgo_c.c     independent C restatement (BFS flood fill) used for
                differential testing and as the fast rules CPU baseline
  go_textbook.py / go_textbook.c
                the rules of Go as a rule book states them (play the stone,
                remove dead opponent groups, suicide test, simple ko, area
                score) -- a third checker that does not know GymGo's
                liberty-count formula; equal to the two restatements on every
                position of BASELINE configs[2] (tests/test_rules_textbook.py)
  nets_ref.py   plain-PyTorch fp32 restatement of the reference networks
                (main.py:72-142), pinned by tests/golden/nets_*.npz
  mcts_ref.py   restatement of the reference MCTS variants
                (main.py:247-368, play.py:104-216), pinned by
                tests/golden/mcts_*.npz
  ref_loader.py imports /root/reference/*.py with gym stubs (only works in
                the build container; used by make_golden.py)
"""
