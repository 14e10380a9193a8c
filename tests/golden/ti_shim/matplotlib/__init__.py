"""Import stub of matplotlib -- TEST INFRASTRUCTURE ONLY (the reference's trainer / controller create
a figure at construction and plot only when densification runs; neither is on the rasteriser path)."""
