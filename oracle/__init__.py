"""ORACLE -- test infrastructure only (see oracle/oracle.c header).  PARITY UNPINNED for the
training arithmetic: the reference holds none and TensorFlow 1.3 cannot run here."""
