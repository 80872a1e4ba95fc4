"""Ready-made formulations: the reference's biped example and the generic
linear system used by the benchmark configurations."""
