"""B200-native theta-l nonhydrostatic dycore step (HOMME hot path) behind the HOMMEXX C ABI."""
