"""CPU oracle — TEST INFRASTRUCTURE (see binding.py).  Not imported by the product package."""
