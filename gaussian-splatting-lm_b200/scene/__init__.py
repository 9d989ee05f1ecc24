"""Only the thin, hot-path part of the reference's `scene` package: the Gaussian parameter
container with its activations and forward-AD plumbing.  Dataset IO, densification, PLY and optimiser setup are
out of scope."""
