"""Import stub of pytorch3d -- TEST INFRASTRUCTURE ONLY (used by the reference's FTGMM diagnostics,
which run every 1234 training iterations and are outside the rasteriser path)."""
