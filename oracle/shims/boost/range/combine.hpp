// Test-infrastructure shim (oracle/_ref build only). Not product code.
