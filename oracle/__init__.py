"""TEST INFRASTRUCTURE ONLY: CPU restatement and compiled builds of the reference LD-score path (the checker, never shipped)."""
