"""TEST INFRASTRUCTURE -- nmigen.hdl subset (see ../__init__.py)."""
