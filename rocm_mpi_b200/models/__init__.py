"""Applications ("model families" of this framework): the diffusion solvers and the p2p smoke test."""
