"""CPU restatement of the reference path — test infrastructure only (see oracle.cpp)."""
