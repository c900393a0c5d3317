"""TEST INFRASTRUCTURE ONLY — CPU/fp32 parity oracle of the reference's hot path.
Never imported by masterthesis_b200/ (the product)."""
