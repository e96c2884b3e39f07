"""Experiment-runner side of the drop-in: the NVML energy-profiler plugin (`nvml_energy`), the compatibility
shims that let the UNCHANGED reference runner start on Linux (`shims/`), and the launcher (`launch`)."""
