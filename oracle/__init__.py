"""CPU oracle — TEST INFRASTRUCTURE ONLY (see oracle/dls_oracle.c). Never imported by the product package."""
