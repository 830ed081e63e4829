"""TEST INFRASTRUCTURE ONLY: CPU restatement of the reference learner path (see learner_oracle.py)."""
