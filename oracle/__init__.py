"""TEST INFRASTRUCTURE ONLY. CPU oracle for the decode hot path (see oracle/oracle.py)."""
