// filled in below (PMJ / HDRI / path-tracer restatement)
