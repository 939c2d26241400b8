"""B200-native (sm_100a) Conformer CTC/attention train step behind the ESPnet2 module APIs."""
__version__ = "0.1.0"
