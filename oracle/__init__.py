"""CPU golden rasteriser -- TEST INFRASTRUCTURE ONLY (see golden_raster.c header)."""
