// empty stand-in for host emulation
