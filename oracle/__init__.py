"""CPU oracle — test infrastructure only.  See oracle/cg_ref.c and oracle/mg_ref.c."""
