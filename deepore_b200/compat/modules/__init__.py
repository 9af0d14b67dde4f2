"""`modules` package name of the reference (src/modules/), resolved to deepore_b200 (see deepore_b200/compat)."""
