"""CPU oracle for the ORB front-end (test infrastructure, never on the product path)."""
