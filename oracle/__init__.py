"""ORACLE — test infrastructure only (see graphmet_oracle.py header). Not a product path."""
