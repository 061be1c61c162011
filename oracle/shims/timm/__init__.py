"""Oracle shim (test infrastructure): minimal stand-in for timm==0.9.16 (requirements/requirements.txt:19)."""
