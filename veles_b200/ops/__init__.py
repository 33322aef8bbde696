"""Hot-path operators: hand-written sm_100a kernels (``kernels``) and their numpy oracles (``oracle``)."""
