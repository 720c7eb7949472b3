"""CPU oracle: test infrastructure only (tests/, __graft_entry__.smoke(), bench.py CPU-baseline legs).
The product package robustpointclouds_b200 never imports this."""
