"""TEST INFRASTRUCTURE ONLY: CPU parity oracles for the paged-KV hot path.

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs.  The product package (light_vllm_b200) never imports this.
"""
