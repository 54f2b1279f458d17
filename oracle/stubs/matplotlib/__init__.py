"""Import stub (see ../README.md)."""
