"""Algorithm use-cases ("models"): matrix factorisation and passive-aggressive."""
