"""TEST INFRASTRUCTURE ONLY — stand-in package so that the UNMODIFIED
/root/reference/pykSpider/kSpider2/ks_pairwise.py (`--estimate-ani` branch, :29-82) runs in this image.
See distance_utils.py."""
