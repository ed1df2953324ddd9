"""tdod: B200-native proposal-and-scoring path of luchsonice/3dod (host side over libtdod.so)."""
