"""Test-infrastructure oracle (CPU restatement of the reference path). Not product code."""
