"""Multi-rank layer: runtime, decomposition, arrays, peer memory, fused exchange plans, collectives."""
