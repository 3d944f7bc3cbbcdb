"""Test infrastructure: CPU oracle for the VISinger synthesis forward (see visinger_oracle.py)."""
