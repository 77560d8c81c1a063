"""Python frontend of the B200 HEAT path: same classes, CLI and YAML semantics as the reference's
cf_cpu/cf/ package, on top of the `cf_c` pybind11 module built from heat_b200/csrc/host."""
