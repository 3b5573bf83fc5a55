"""Stand-in for the two Biopython calls share/palace/scripts/get_ref_by_index.py makes (Biopython is not in
this image).  Test infrastructure only: used by tests/golden/make_golden_f1.py to run the UNMODIFIED reference
script; never imported by the product."""
