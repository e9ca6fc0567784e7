"""kbmod_b200: B200-native (sm_100a) shift-and-stack trajectory search behind the ``kbmod.search`` API."""
