"""TEST INFRASTRUCTURE: CPU restatement of the reference path (see oracle/model.py). Not imported by the product."""
