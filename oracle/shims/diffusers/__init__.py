"""Restatement of the tiny slice of diffusers==0.27.2 (requirements.txt:4 of the
reference; not vendored, not installable here) that the flow estimator uses.
Test infrastructure only: used to import the *real* reference modules from
/root/reference inside the build container when generating golden vectors."""
