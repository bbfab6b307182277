"""TEST INFRASTRUCTURE ONLY -- see oracle/seqalign_oracle.c.  Imported by tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs; never by the product
package sequencealigner_b200."""
