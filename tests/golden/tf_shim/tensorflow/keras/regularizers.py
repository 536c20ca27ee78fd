def l2(l2=0.01):
    """Declared by the reference but never applied: a no-op marker here."""
    return ("l2", l2)
