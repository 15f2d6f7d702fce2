"""Import stub (tests only): the entropy coder used by the reference's compress()/decompress(); not on the render path."""
