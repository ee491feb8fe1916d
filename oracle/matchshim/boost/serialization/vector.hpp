// oracle/matchshim: empty stand-in for <boost/serialization/vector.hpp> (test infrastructure); names come from facade.h
