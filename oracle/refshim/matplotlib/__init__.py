"""Empty stand-in for matplotlib (imported by utils/logger.py:31). TEST INFRASTRUCTURE."""
