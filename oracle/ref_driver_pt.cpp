// filled in below
