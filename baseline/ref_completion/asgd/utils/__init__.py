# ABSENT in the reference; supplied so the shipped, byte-identical files import.
