"""Drop-ins for the reference's ``unziplora_unet`` package (hot-path modules only)."""
