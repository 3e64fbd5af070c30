"""TEST INFRASTRUCTURE ONLY — import stubs that let the UNMODIFIED reference code under
/root/reference run on CPU torch inside the build container (no Isaac Sim, no skrl).

Nothing in here is shipped, nothing in here runs on the GPU box (``/root/reference`` does not
exist there).  It is used by ``oracle/gen_golden.py`` to (a) pin the numpy/torch restatement in
``oracle/`` against the reference's own code and (b) write the golden vectors under
``tests/golden/``.  Recipe: SURVEY.md Appendix C.
"""
