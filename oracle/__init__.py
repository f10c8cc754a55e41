"""CPU oracle for the NDS + crowding path.  Test infrastructure only -- see nds_oracle.py."""
