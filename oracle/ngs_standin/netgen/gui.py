"""No GUI on this box: importing is a no-op."""
