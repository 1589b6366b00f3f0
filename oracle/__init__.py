"""Test infrastructure only: CPU restatement of the reference path (mt_oracle) and the recipe that
builds the reference's own kernels as a checker (build_ref.sh -> oracle/_ref/)."""
