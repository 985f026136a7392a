"""CPU oracle for the object-centric PPO learner path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oc_cleanrl_b200/`` may import this
package.  Allowed callers: ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` -- and there only
as the checker / the CPU arm, never as the thing shipped.

Parity status: PINNED against the reference's own ``PPObj`` class imported
from ``/root/reference/cleanrl/architectures/ppo.py`` in the build container
(``tests/test_oracle_cpu.py::test_oracle_model_equals_reference_model``; golden vectors written by
``tests/golden/make_golden.py`` and committed under ``tests/golden/``).  The
reference's learner loop is inline ``__main__`` code that cannot be imported
(its module-level imports need gymnasium / stable_baselines3 / ocatari_wrappers,
all absent), so GAE / loss / clip / Adam are restated operator by operator with
the same torch calls; the reference ships no golden vectors for them
(SURVEY.md section 8c).
"""
