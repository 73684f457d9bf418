"""TEST INFRASTRUCTURE -- nmigen.lib subset (see ../__init__.py)."""
