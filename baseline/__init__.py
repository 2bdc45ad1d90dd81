"""The reference arm of the benchmark: the UNMODIFIED reference rasterizer extension, built from
/root/reference/submodules/diff-gaussian-rasterization by baseline/build_ref.sh into the git-ignored
baseline/_ref/ (see baseline/reference.py).  Nothing here is product code; only bench.py --impl
reference and the parity tests touch it."""
