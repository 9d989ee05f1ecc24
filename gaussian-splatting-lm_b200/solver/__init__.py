"""Solver-side interface of the LM hot path (reference package `solver`): generalized vectors, the matvec
callables and CG / CGLS."""
