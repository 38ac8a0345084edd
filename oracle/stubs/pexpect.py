"""Empty stub: the reference's server_manager.py:3 imports pexpect for SSH control,
which is outside the routing hot path (TEST INFRASTRUCTURE ONLY)."""
