"""Drop-ins for the reference's ``animatediff`` package (hot-path modules only)."""
