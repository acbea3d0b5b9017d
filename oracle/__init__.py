"""CPU oracle package — TEST INFRASTRUCTURE ONLY (see oracle/README.md). Never imported by mpcb200."""
